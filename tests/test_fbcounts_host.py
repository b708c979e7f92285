"""Host side of getCounts (historian_b200/fbcounts.py): the alignment columns ForwardMatrix::getAlignmentColumn
(forward.cpp:940-973) builds one cell at a time, assembled as one weighted batch.  No GPU needed."""
import numpy as np

from historian_b200 import fbcounts as F

GAP, WILD = ord("-"), ord("*")


def test_leaf_columns_follow_profile_align_column():
    cols = F.leaf_columns("acg", 1, 3)
    assert cols.shape == (5, 3)                       # START, three residues, END
    assert bytes(cols[:, 1]) == b"-acg-"
    assert np.all(cols[:, [0, 2]] == GAP)
    assert F.leaf_columns(b"", 0, 3).shape == (2, 3)  # empty sequence: START and END only


def test_column_batch_classes_order_and_weights():
    x, y = F.leaf_columns("ac", 0, 3), F.leaf_columns("gt", 1, 3)
    w_x = np.array([[0, 0.25, 0, 0], [0, 0.5, 0.125, 0]])     # insert / delete weights per x state
    w_y = np.array([[0, 0, 0, 0], [0, 0, 0.75, 0]])
    cells = np.array([[1, 1, 0], [2, 2, 0]], np.int32)
    cols, w = F.column_batch(x, y, 2, w_x, w_y, cells, np.array([0.3, 0.6]))
    got = [(bytes(c).decode(), float(v)) for c, v in zip(cols, w)]
    assert got == [("a--", 0.25),                 # x insert: the child's own column, no parent
                   ("a-*", 0.5), ("c-*", 0.125),  # x deleted: wildcard parent
                   ("-t*", 0.75),                 # y deleted
                   ("ag*", 0.3), ("ct*", 0.6)]    # matches: both children and the parent


def test_column_batch_skips_empty_columns_and_merges_tree_rows():
    # internal profiles: columns span several tree rows; a null state's column is empty and never emitted
    x = np.array([list(b"-----"), list(b"a-c--"), list(b"-----"), list(b"-----")], np.uint8)
    y = np.array([list(b"-----"), list(b"---g-"), list(b"-----")], np.uint8)
    w_x = np.array([[0, 0.5, 0.5, 0], [0, 0.25, 0.25, 0]])
    w_y = np.zeros((2, 3))
    cols, w = F.column_batch(x, y, 4, w_x, w_y, np.array([[1, 1, 0]], np.int32), np.array([1.0]))
    got = [(bytes(c).decode(), float(v)) for c, v in zip(cols, w)]
    assert got == [("a-c--", 0.5), ("a-c-*", 0.25), ("a-cg*", 1.0)]


def test_prob_model_waits_order():
    class P:
        def __init__(self, t, i, d):
            self.t, self.insWait, self.delWait = t, i, d
    assert list(F.prob_model_waits(P(1, 2, 3), P(4, 5, 6))) == [1, 4, 2, 3, 5, 6]


def test_count_propagation_fixtures_state_the_open_target():
    """tests/golden/fbcounts_prop_*.npz (tools/make_golden_fbcounts.py, REF_DUMP_COUNTS=2): the reference's getCounts at the
    root of a reconstruction made WITH CountSubstEvents | CountIndelEvents, i.e. with counts carried by the transitions of
    the internal profiles (forward.cpp:753-790) -- the part of row B3 the library does not build yet (DESIGN.md section 7,
    item 1).  Nothing on the device is compared here; the test pins what the fixtures hold so that the target cannot rot:
    the children's transitions carry counts, and the root's totals contain the whole subtree's events, not only the root
    merge's (the count-free fixtures of the same runs)."""
    import glob
    import os
    import helpers as H
    props = sorted(glob.glob(os.path.join(H.GOLDEN_DIR, "fbcounts_prop_*.npz")))
    assert len(props) == 4
    for path in props:
        d = H.load_golden(path)
        base = H.load_golden(path.replace("fbcounts_prop_", "fbcounts_"))
        root = int(d["root"][0])
        pfx = "n%d." % root
        assert np.array_equal(d[pfx + "x.trans_src"], base[pfx + "x.trans_src"])          # same reconstruction
        assert d[pfx + "lp_end"][0] == base[pfx + "lp_end"][0]
        for side in ("x.", "y."):
            nt = int(d[pfx + side + "dims"][1])
            ti, ts = d[pfx + side + "trans_indel"], d[pfx + side + "trans_sub_count"]
            assert ti.shape == (nt, 7) and ts.shape[0] == nt
            assert np.all(ti[:, :6] >= 0) and ti[:, :6].sum() > 0                            # transitions carry indel counts
            assert np.all(np.isfinite(ts))     # substitution counts only where a path eliminated an inserted column
        pi, bi = d[pfx + "fbc.indel"], base[pfx + "fbc.indel"]
        assert pi[6] == bi[6]                                                                # lp = the root's Forward likelihood
        assert np.all(pi[:6] >= bi[:6] - 1e-9) and pi[4] > bi[4] and pi[5] > bi[5]           # wait times of every branch, not two
        assert d[pfx + "fbc.sub_count"].sum() >= base[pfx + "fbc.sub_count"].sum() - 1e-9


def test_oracle_indel_counts_with_transition_counts_equal_the_reference():
    """oracle_fb_indel_counts -- the indel half of getCounts with the children's per-transition counts (forward.cpp:579-583,
    1183-1214) -- against the reference on all eight tree fixtures: the count-free runs (transitions carry nothing) and the
    runs made with the Count* strategy bits, whose per-transition counts the fixtures hold.  The exact-arithmetic builds
    must agree to 1e-9; the table builds to the table's error.  This pins the oracle for the open part of B3's indel side."""
    import glob
    import os
    import helpers as H
    import oracle_lib as O
    paths = sorted(glob.glob(os.path.join(H.GOLDEN_DIR, "fbcounts_*tree*.npz")))
    assert len(paths) == 8
    for path in paths:
        d = H.load_golden(path)
        root = int(d["root"][0])
        pfx = "n%d." % root
        x, y, hmm, env = H.node_inputs(d, root)
        prop = "_prop_" in os.path.basename(path)
        xt = d[pfx + "x.trans_indel"][:, :6] if prop else None
        yt = d[pfx + "y.trans_indel"][:, :6] if prop else None
        exact = path.endswith("_exact.npz")
        O.set_exact_lse(exact)
        try:
            got = O.fb_indel_counts(x, y, hmm, env, xt, yt, d[pfx + "fbc.waits"])
        finally:
            O.set_exact_lse(False)
        want = d[pfx + "fbc.indel"]
        tol = 1e-9 if exact else 2e-3
        assert abs(got[6] - want[6]) <= 1e-9 * abs(want[6]), path
        assert np.allclose(got[:6], want[:6], rtol=tol, atol=tol), (os.path.basename(path), got, want)


def test_oracle_counts_through_make_profile_equal_the_reference():
    """oracle_eliminate_indel_counts -- makeProfile's count propagation (forward.cpp:753-790), indel part -- against the
    reference on every non-root merge of the two tree families (tests/golden/elimcounts_*_best.npz: best-trace profiles
    built with the Count* strategy bits): cherries (children without counts) and internal merges whose children's
    transitions already carry counts.  The parent's first n transitions are the effective transitions in elimination order
    (addReadyStates appends its own, count-free ones)."""
    import glob
    import os
    import helpers as H
    import oracle_lib as O
    paths = sorted(glob.glob(os.path.join(H.GOLDEN_DIR, "elimcounts_*_best.npz")))
    assert len(paths) == 2
    checked = with_child_counts = 0
    for path in paths:
        d = H.load_golden(path)
        nodes = [int(v) for v in d["dumped_nodes"]]
        for node in nodes[:-1]:   # the root's profile is built without counts (recon.cpp:1009)
            pfx = "n%d." % node
            x, y, hmm, env = H.node_inputs(d, node)
            xt, yt = d[pfx + "x.trans_indel"][:, :6], d[pfx + "y.trans_indel"][:, :6]
            got = O.eliminate_indel_counts(x, y, hmm, env, d[pfx + "best_trace"], int(d["strategy"][0]), xt, yt, d[pfx + "fbc.waits"])
            want = d[pfx + "parent.trans_indel"]
            n = len(got)
            assert n <= len(want) and np.all(want[n:, :6] == 0), (path, node)     # the ready transitions carry nothing
            assert np.allclose(got, want[:n, :6], rtol=1e-9, atol=1e-12), (os.path.basename(path), node)
            assert want[:n, :6].sum() > 0
            checked += 1
            with_child_counts += bool(xt.sum() + yt.sum() > 0)
    assert checked == 6 and with_child_counts >= 2
