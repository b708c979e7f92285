"""BackwardMatrix::getCounts (forward.cpp:1183-1214) on the device (hist_fb_counts + a weighted
hist_sumprod_columns batch, driven by historian_b200/fbcounts.py) against the REAL reference's getCounts() on
two-leaf families: tests/golden/fbcounts_*.npz, made by tools/make_golden_fbcounts.py from
oracle/ref_dump_fbcounts.cpp.  The fixtures hold only sequences, branch lengths, the band and the reference's
outputs; model, PairHMM, leaf profiles, envelope and eigen-system are rebuilt here through the product path.

Tolerance.  The reference adds log-probabilities with a lookup-table log-sum-exp (<= 4.6e-5 per binary op,
logsumexp.h:47) in both fills; the device fills are exact-sum FP32-mantissa + exponent.
Posterior weights therefore differ by a few 1e-4 relative; against the reference AS SHIPPED counts are compared at
2e-3 of the largest count of their block, the Forward likelihood at 5e-3 + 1e-4 relative (the bound test_fill_gpu.py
uses against goldens).

The arbiter is the reference itself in exact arithmetic: the fixtures fbcounts_*_exact.npz come from the same unmodified
sources compiled with the reference's own -DLOG_SUM_EXP_SLOW switch (logsumexp.h:12-15; oracle/Makefile ref_exact).
Against those EVERY count is compared on its own at 1e-4 relative (BASELINE north_star's bound), small off-diagonal
substitution counts and rare indel events included, and the likelihood at 2e-6 relative + 2e-4."""
import glob
import os

import numpy as np
import pytest

import helpers as H
from historian_b200 import capi, descs as D, fbcounts, model as M

pytestmark = pytest.mark.gpu
ALL = sorted(f for f in glob.glob(os.path.join(H.GOLDEN_DIR, "fbcounts_*.npz")) if "_prop_" not in os.path.basename(f))   # fbcounts_prop_*: tests/test_fbcounts_host.py
FIX = [f for f in ALL if "_tree" not in os.path.basename(f)]         # two leaf profiles
IDS = [os.path.basename(f)[9:-4] for f in FIX]
TREES = [f for f in ALL if "_tree" in os.path.basename(f)]           # root merge of a progressive reconstruction
TREE_IDS = [os.path.basename(f)[9:-4] for f in TREES]


@pytest.fixture(scope="module")
def ctx():
    c = capi.Context(0)
    yield c
    c.close()


def rate_model(spec):
    f = spec.split(":")
    m = M.RateModel.preset(f[0])
    return m.gamma(int(f[1]), float(f[2])) if len(f) == 3 else m


def run_case(ctx, d):
    model = rate_model(bytes(d["preset"]).decode())
    x, y = bytes(d["x_seq"]).decode(), bytes(d["y_seq"]).decode()
    tl, tr = (float(v) for v in d["branch"])
    band = int(d["band"][0])
    env = None
    if band >= 0:
        env = D.Envelope.from_guide_rows(d["x_row"] != ord("-"), d["y_row"] != ord("-"), band)
    dev = ctx.upload_sequences([x, y], model.alphabet, components=model.components)
    lp, rp = M.ProbModel(model, tl), M.ProbModel(model, tr)
    hmm = M.pair_hmm(model, tl, tr)
    pair = ctx.prepare(dev[0], dev[1], hmm, env)
    lp_end = pair.forward()
    pair.backward()
    sp = ctx.sumprod(D.SumProdModel.from_rate_model(model), D.TreeTopology([2, 2, -1], [tl, tr, 0.0]))
    waits = fbcounts.prob_model_waits(lp, rp)
    counts = fbcounts.fb_counts(pair, sp, waits, fbcounts.leaf_columns(x, 0, 3), fbcounts.leaf_columns(y, 1, 3), 2, lp_end)
    root, sub = sp.transform_counts(counts)
    raw = pair.fb_counts(waits)
    sp.close()
    return lp_end, waits, counts, root, sub, raw


def near(a, b, rel):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return bool(np.all(np.abs(a - b) <= rel * max(1e-12, np.abs(b).max())))


def is_exact(path):
    return path.endswith("_exact.npz")


def check_counts(path, lp_end, ref_lp, counts, root, sub, ref_indel, ref_root, ref_sub):
    if is_exact(path):
        # entry by entry against the reference in exact arithmetic
        assert abs(lp_end - ref_lp) <= 2e-6 * abs(ref_lp) + 2e-4, (lp_end, ref_lp)
        for what, got, want in (("indel", counts[0:6], ref_indel[0:6]), ("root", root, ref_root), ("sub", sub, ref_sub)):
            got, want = np.asarray(got, float), np.asarray(want, float)
            err = np.abs(got - want) - (1e-4 * np.abs(want) + 1e-9)
            assert np.all(err <= 0), (what, float(err.max()), got.ravel()[np.argmax(err)], want.ravel()[np.argmax(err)])
        return
    assert abs(lp_end - ref_lp) <= 5e-3 + 1e-4 * abs(lp_end)
    assert near(counts[0:4], ref_indel[0:4], 2e-3), (counts[:6], ref_indel[:6])      # ins, del, insExt, delExt
    assert near(counts[4:6], ref_indel[4:6], 2e-3), (counts[:6], ref_indel[:6])      # insTime, delTime
    assert near(root, ref_root, 2e-3)
    assert near(sub, ref_sub, 2e-3)


@pytest.mark.parametrize("path", FIX, ids=IDS)
def test_counts_match_reference_getcounts(ctx, path):
    d = H.load_golden(path)
    lp_end, waits, counts, root, sub, _ = run_case(ctx, d)
    assert np.allclose(waits, d["waits"], rtol=1e-12, atol=0)
    assert counts[6] == lp_end
    check_counts(path, lp_end, d["lp_end"][0], counts, root, sub, d["indel"], d["root_count"], d["sub_count"])


@pytest.mark.parametrize("path", FIX, ids=IDS)
def test_posterior_weights_are_consistent(ctx, path):
    """Size-independent properties: every x residue is absorbed exactly once along any path, so its posterior
    weights over match, deletion and insertion columns sum to one (up to the insert-cache quirk, which only
    zeroes weights; 1e-3 covers the FP32 mantissa of the stored cells); the run is bit-reproducible."""
    d = H.load_golden(path)
    _, waits, counts, _, _, (indel, w_x, w_y, cells, cell_w) = run_case(ctx, d)
    _, _, counts2, _, _, raw2 = run_case(ctx, d)
    assert H.same(counts, counts2) and all(H.same(a, b) for a, b in zip((indel, w_x, w_y, cells, cell_w), raw2))
    xn, yn = w_x.shape[1], w_y.shape[1]
    mx, my = np.zeros(xn), np.zeros(yn)
    np.add.at(mx, cells[:, 0], cell_w)
    np.add.at(my, cells[:, 1], cell_w)
    banded = int(d["band"][0]) >= 0
    for w, m, n, side in ((w_x, mx, xn, "x"), (w_y, my, yn, "y")):
        tot = (w[0] + w[1] + m)[1:n - 1]
        kept = w[0][1:n - 1] > 0          # rows whose insert-column cache the reference did not leave empty
        assert np.all(tot <= 1 + 1e-3)
        assert np.all(np.abs(tot[kept] - 1) <= 1e-3), tot
        if side == "y" or not banded:     # xNearStart[0] puts all of row 0 inside the envelope (forward.cpp:58): the
            assert not kept.any()         # y cache is always filled from xpos == 0; without a band so is the x cache
        else:
            assert kept.sum() >= (n - 2) // 2
    assert np.all(cells[:, 2] == D.IMM) and np.all(np.diff(cells[:, 0] * 100000 + cells[:, 1]) > 0)


@pytest.mark.parametrize("path", TREES, ids=TREE_IDS)
def test_root_merge_of_a_progressive_reconstruction(ctx, path):
    """getCounts at the root of a 4- and a 6-leaf reconstruction: DAG profiles with null states, alignment
    columns spanning several tree rows (Profile::alignColumn, profile.cpp:100-110, dumped by the reference), a
    gamma mixture.  Profiles, PairHMM and envelope come from the reference's dump, as in test_fill_gpu.py."""
    d = H.load_golden(path)
    root = int(d["root"][0])
    pfx = "n%d." % root
    x, y, hmm, env = H.node_inputs(d, root)
    model = rate_model(bytes(d["preset"]).decode())
    pair = ctx.prepare(ctx.upload(x), ctx.upload(y), hmm, env)
    lp_end = pair.forward()
    pair.backward()
    sp = ctx.sumprod(D.SumProdModel.from_rate_model(model), D.TreeTopology(d["tree.parent"], d["tree.branch_length"]))
    counts = fbcounts.fb_counts(pair, sp, d[pfx + "fbc.waits"], d[pfx + "x.align_col"], d[pfx + "y.align_col"], root, lp_end)
    rc, sub = sp.transform_counts(counts)
    sp.close()
    check_counts(path, lp_end, d[pfx + "lp_end"][0], counts, rc, sub, d[pfx + "fbc.indel"], d[pfx + "fbc.root_count"], d[pfx + "fbc.sub_count"])


def test_full_size_pair_properties(ctx):
    """BASELINE-sized merge (two 1000-residue LG leaves, guide band 20, visible indel rates): properties that do
    not need the reference - posterior mass per residue, the expected number of root residues, the closed-form
    link between indel events and waiting times, and bit-reproducibility."""
    from historian_b200 import simulator as S
    lg = M.RateModel.preset("lg")
    lg.insRate = lg.delRate = 0.05
    t = 0.2
    tree = S.Tree.balanced(2, t)
    gapped, _ = S.simulate_family(lg, tree, 1000, 11)
    _, guide = S.leaf_guide(tree, gapped)
    gap = ord("-")
    seqs = [bytes(r[r != gap]) for r in guide[:2]]
    env = D.Envelope.from_guide_rows(guide[0] != gap, guide[1] != gap, 20)
    x, y = ctx.upload_sequences(seqs, lg.alphabet, lg.components)
    pair = ctx.prepare(x, y, M.pair_hmm(lg, t, t), env)
    lp_end = pair.forward()
    pair.backward()
    lp, rp = M.ProbModel(lg, t), M.ProbModel(lg, t)
    waits = fbcounts.prob_model_waits(lp, rp)
    indel, w_x, w_y, cells, cell_w = pair.fb_counts(waits)
    again = pair.fb_counts(waits)
    assert all(H.same(a, b) for a, b in zip((indel, w_x, w_y, cells, cell_w), again))
    xn, yn = w_x.shape[1], w_y.shape[1]
    assert (xn, yn) == (len(seqs[0]) + 2, len(seqs[1]) + 2)
    mx, my = np.zeros(xn), np.zeros(yn)
    np.add.at(mx, cells[:, 0], cell_w)
    np.add.at(my, cells[:, 1], cell_w)
    kept = w_x[0] > 0
    # (the last residues of x and y can only be matched to each other: neither profile is 'ready' in its last state,
    # forward.cpp:97, so row L has no insert or delete weight at all)
    assert mx[xn - 2] == 1.0 and kept[80:xn - 2].all() and not (w_y[0] > 0).any()      # the insert-cache quirk: x rows next to the start, all of y
    assert np.all(np.abs((w_x[0] + w_x[1] + mx)[kept] - 1) <= 2e-3)
    assert np.all((w_y[1] + my)[1:-1] <= 1 + 2e-3)
    # every event count is a posterior expectation of a non-negative quantity
    assert np.all(indel >= 0) and indel[0] + indel[1] > 0.5
    # waiting times (forward.cpp:590-640): a match adds at most both branch lengths (exactly both when it follows a
    # match), a deletion at most one branch, an opening event one decay wait
    matches, dels_x, dels_y = cell_w.sum(), w_x[1].sum(), w_y[1].sum()
    lo = 0.9 * matches * 2 * t
    hi = (matches + dels_x + dels_y + 50) * 2 * t + (indel[0] + indel[1]) * max(waits[2:])
    assert lo <= indel[4] <= hi and lo <= indel[5] <= hi
    sp = ctx.sumprod(D.SumProdModel.from_rate_model(lg), D.TreeTopology([2, 2, -1], [t, t, 0.0]))
    counts = fbcounts.fb_counts(pair, sp, waits, fbcounts.leaf_columns(seqs[0], 0, 3), fbcounts.leaf_columns(seqs[1], 1, 3), 2, lp_end)
    root, sub = sp.transform_counts(counts)
    sp.close()
    # one root residue per weighted column: matches, deletions, and the x inserts whose cache is not empty
    assert abs(root.sum() - (matches + dels_x + dels_y + w_x[0].sum())) <= 1e-6 * root.sum()
    assert np.all(root >= -1e-9) and np.all(np.diagonal(sub, axis1=1, axis2=2) >= -1e-9)
