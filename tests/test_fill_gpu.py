"""GPU parity tests of the DPMatrix constructor tables and the Forward/Backward fills,
through the C ABI, against (a) the committed golden fixtures produced by the real
reference and (b) the CPU oracle on seeded synthetic inputs.

Tolerances (stated once, used everywhere below)
  * integer band/envelope bookkeeping: bit-exact.
  * constructor tables (leftMultiply, insx, rootsubx): the GPU computes them exactly in
    FP64 probability space; the reference uses A table log-sum-exps per entry, each up to
    4.54e-5 low (logsumexp.h:47 truncation) -> |diff| <= A * 4.6e-5 against the golden,
    <= 1e-9 against the oracle run with the reference's exact (*_slow) log-sum-exp.
  * Forward/Backward cells and log-likelihoods: north_star's bound is 1e-4 RELATIVE.
    Against the golden (table LSE): |diff| <= 1e-4*|ref| + 5e-3 (the table error
    accumulates along the path).  Against the exact-LSE oracle: |diff| <= 2e-6*|ref| + 2e-4
    (FP32 linear-space mantissas, 2^-24 per operation).
"""
import os

import numpy as np
import pytest

import helpers as H
import oracle_lib as O
import synth
from historian_b200 import descs as D

pytestmark = pytest.mark.gpu

CASES = H.golden_cases()
REL_GOLD, ABS_GOLD = 1e-4, 5e-3
REL_EXACT, ABS_EXACT = 2e-6, 2e-4


@pytest.fixture(scope="module")
def goldens():
    return {f: H.load_golden(f) for f in H.golden_files()}


def prepare(ctx, x, y, hmm, env):
    dx, dy = ctx.upload(x), ctx.upload(y)
    return ctx.prepare(dx, dy, hmm, env), dx, dy


def check_against_oracle(ctx, x, y, hmm, env, what, backward=True):
    pair, dx, dy = prepare(ctx, x, y, hmm, env)
    # --- bit-exact integers
    ot = O.pair_tables(x, y, hmm, env)
    ge = pair.envelope(full=True)
    for k in ("x_near_start", "y_near_end", "in_envelope"):
        assert H.same(ge[k], ot[k]), (what, k)
    if env.max_distance >= 0:
        assert H.same(ge["x_key"], ot["x_key"]) and H.same(ge["y_key"], ot["y_key"])
    assert pair.n_cells == ot["n_cells"], what
    # --- exact-LSE oracle
    O.set_exact_lse(True)
    try:
        et = O.pair_tables(x, y, hmm, env)
        efwd, elp_end = O.forward(x, y, hmm, env)
        ebwd, elp_start = O.backward(x, y, hmm, env) if backward else (None, None)
    finally:
        O.set_exact_lse(False)
    gt = pair.tables()
    # one-component codon alphabets take the tensor-core constructor (hist_codon.cu: 3xTF32 products, FP32 accumulation)
    tensor = x.components == 1 and 32 < x.alph_size <= 64
    for k in ("subx", "suby", "insx", "insy", "rootsubx", "rootsuby"):
        H.assert_log_close(gt[k][..., None], et[k][..., None], 1e-12, 5e-6 if tensor else 1e-9, what=what + ":" + k)
    lp_end = pair.forward()
    gfwd = pair.cells_dense(0)
    H.assert_log_close(gfwd, efwd, REL_EXACT, ABS_EXACT, what=what + ":fwd")
    if np.isfinite(elp_end):
        assert abs(lp_end - elp_end) <= REL_EXACT * abs(elp_end) + ABS_EXACT, (what, lp_end, elp_end)
    else:
        assert lp_end == elp_end
    if backward:
        lp_start = pair.backward()
        gbwd = pair.cells_dense(1)
        H.assert_log_close(gbwd, ebwd, REL_EXACT, ABS_EXACT, what=what + ":bwd")
        if np.isfinite(elp_start):
            assert abs(lp_start - elp_start) <= REL_EXACT * abs(elp_start) + ABS_EXACT
            # Forward == Backward (the reference checks this at 1e-2 relative, forward.cpp:1091)
            assert abs(lp_start - lp_end) <= 1e-5 * abs(lp_end) + 1e-3
    # --- and the reference's own table arithmetic
    tfwd, tlp_end = O.forward(x, y, hmm, env)
    H.assert_log_close(gfwd, tfwd, REL_GOLD, ABS_GOLD, what=what + ":fwd-vs-table")
    return pair


@pytest.mark.parametrize("cid,path,node", CASES, ids=[c[0] for c in CASES])
def test_golden_fixture(ctx, goldens, cid, path, node):
    """GPU vs the REAL reference's outputs (golden fixtures)."""
    d = goldens[path]
    pfx = "n%d." % node
    x, y, hmm, env = H.node_inputs(d, node)
    pair, dx, dy = prepare(ctx, x, y, hmm, env)
    ge = pair.envelope(full=True)
    assert H.same(ge["x_near_start"], d[pfx + "x_near_start"])
    assert H.same(ge["y_near_end"], d[pfx + "y_near_end"])
    assert H.same(ge["in_envelope"], d[pfx + "in_envelope"])
    assert pair.n_cells == int(d[pfx + "in_envelope"].sum())
    A = x.alph_size
    gt = pair.tables()
    for k in ("subx", "suby", "insx", "insy", "rootsubx", "rootsuby"):
        H.assert_log_close(gt[k][..., None], d[pfx + k][..., None], 0.0, (A * x.components + 1) * 4.6e-5, what=cid + ":" + k)
    lp_end = pair.forward()
    ref_end = float(d[pfx + "lp_end"][0])
    assert abs(lp_end - ref_end) <= REL_GOLD * abs(ref_end) + ABS_GOLD, (lp_end, ref_end)
    H.assert_log_close(pair.cells_dense(0), d[pfx + "fwd"], REL_GOLD, ABS_GOLD, what=cid + ":fwd")
    lp_start = pair.backward()
    ref_start = float(d[pfx + "lp_start"][0])
    assert abs(lp_start - ref_start) <= REL_GOLD * abs(ref_start) + ABS_GOLD
    H.assert_log_close(pair.cells_dense(1), d[pfx + "bwd"], REL_GOLD, ABS_GOLD, what=cid + ":bwd")
    # cell order of hist_cells_fetch == DPMatrix::write order (row-major over the envelope)
    lp, xp, yp = pair.cells(0)
    ii, jj = np.nonzero(d[pfx + "in_envelope"])
    assert H.same(xp, ii.astype(np.int32)) and H.same(yp, jj.astype(np.int32))


@pytest.mark.parametrize("cid,path,node", CASES[:6], ids=[c[0] for c in CASES[:6]])
def test_golden_inputs_vs_exact_oracle(ctx, goldens, cid, path, node):
    x, y, hmm, env = H.node_inputs(goldens[path], node)
    check_against_oracle(ctx, x, y, hmm, env, cid)


@pytest.mark.parametrize("seed", range(8))
def test_synthetic_dag_profiles(ctx, seed):
    """Random state DAGs with null/wait/ready states, skip transitions, mixtures, with and
    without band, monotone and non-monotone closest-leaf coordinates."""
    rng = np.random.default_rng(1000 + seed)
    A = [4, 20, 4, 20, 7, 20, 4, 61][seed]
    C = [1, 1, 2, 1, 3, 4, 1, 1][seed]
    nx, ny = int(rng.integers(5, 70)), int(rng.integers(5, 70))
    x = synth.random_profile(rng, nx, A, C, one_hot=seed % 2 == 0, jitter_leaf=seed >= 4)
    y = synth.random_profile(rng, ny, A, C, one_hot=seed % 2 == 0, jitter_leaf=seed >= 4)
    hmm = synth.random_hmm(rng, A, C)
    env = D.Envelope(-1) if seed % 3 == 0 else synth.random_envelope(rng, x, y, int(rng.integers(0, 8)))
    check_against_oracle(ctx, x, y, hmm, env, "synth%d" % seed)


def test_leaf_pair_banded_1000(ctx):
    """Leaf x leaf merge of ~1000-residue amino sequences inside the default band (20): the
    BASELINE config-3 cherry shape, at full size (the oracle needs a few seconds)."""
    rng = np.random.default_rng(7)
    alphabet = "arndcqeghilkmfpstwyv"
    x, y, (r1, r2) = synth.leaf_pair(rng, alphabet, 1000, 990)
    hmm = synth.random_hmm(rng, 20, 1, tl=0.05, tr=0.05)
    env = D.Envelope.from_guide_rows(r1, r2, 20)
    pair = check_against_oracle(ctx, x, y, hmm, env, "leaf1000")
    assert pair.info.max_diag_width <= 48


def test_leaf_pair_unbanded_300(ctx):
    rng = np.random.default_rng(8)
    alphabet = "acgt"
    x, y, _ = synth.leaf_pair(rng, alphabet, 300, 280)
    hmm = synth.random_hmm(rng, 4, 2)
    x = D.Profile(x.is_emit, x.trans_src, x.trans_dest, x.trans_lp, np.repeat(x.lp_absorb, 2, axis=1), x.closest_leaf_pos)
    y = D.Profile(y.is_emit, y.trans_src, y.trans_dest, y.trans_lp, np.repeat(y.lp_absorb, 2, axis=1), y.closest_leaf_pos)
    check_against_oracle(ctx, x, y, hmm, D.Envelope(-1), "leaf300-noband")


def test_edge_cases_empty_and_tiny(ctx):
    """Zero-length sequences (xEmpty / yEmpty, forward.cpp:16-17,97; the reference's testzerolen)
    and single-residue sequences."""
    rng = np.random.default_rng(3)
    alphabet = "acgt"
    hmm = synth.random_hmm(rng, 4, 1)
    for sx, sy in (("", ""), ("", "acg"), ("a", ""), ("a", "c"), ("ac", "a")):
        x, y = D.Profile.from_sequence(sx, alphabet), D.Profile.from_sequence(sy, alphabet)
        check_against_oracle(ctx, x, y, hmm, D.Envelope(-1), "tiny-%s-%s" % (sx, sy))


def test_zero_likelihood_band_is_not_an_error(ctx):
    """lpEnd = -inf must come back as a value (the caller widens the band, recon.cpp:956-975)."""
    rng = np.random.default_rng(5)
    alphabet = "acgt"
    x, y = D.Profile.from_sequence("a" * 30, alphabet), D.Profile.from_sequence("c" * 3, alphabet)
    hmm = synth.random_hmm(rng, 4, 1)
    # guide: all of x first, then y: band 0 leaves no path from start to end
    r1 = np.array([1] * 30 + [0] * 3, bool)
    r2 = np.array([0] * 30 + [1] * 3, bool)
    env = D.Envelope.from_guide_rows(r1, r2, 0)
    ofwd, olp = O.forward(x, y, hmm, env)
    pair, dx, dy = prepare(ctx, x, y, hmm, env)
    lp = pair.forward()
    assert lp == olp or abs(lp - olp) < 1e-3 + 1e-4 * abs(olp)


def test_batch_fill_matches_single(ctx):
    """hist_fill_batch (one launch for all sibling merges of a level) == per-pair fills."""
    rng = np.random.default_rng(11)
    alphabet = "arndcqeghilkmfpstwyv"
    hmm = synth.random_hmm(rng, 20, 1, tl=0.1, tr=0.2)
    keep, singles, xs, ys, envs = [], [], [], [], []
    for k in range(9):
        x, y, (r1, r2) = synth.leaf_pair(rng, alphabet, 60 + 7 * k, 50 + 11 * k)
        env = D.Envelope.from_guide_rows(r1, r2, 5 + k) if k != 4 else None
        q, ex, ey = prepare(ctx, x, y, hmm, env if env is not None else D.Envelope(-1))
        xs.append(ctx.upload(x)); ys.append(ctx.upload(y)); envs.append(env)
        keep += [ex, ey]; singles.append(q)
    # hist_pair_prepare_batch: one transfer + one launch per table kernel for the whole level
    pairs = ctx.prepare_batch(xs, ys, hmm, envs, backward=True)
    for p, q in zip(pairs, singles):
        assert p.n_cells == q.n_cells
        tp, tq = p.tables(), q.tables()
        for key in tp:
            assert H.same(tp[key], tq[key]), key
    lp_end, lp_start = ctx.fill_batch(pairs, backward=True)
    for k, q in enumerate(singles):
        assert q.forward() == lp_end[k]
        assert q.backward() == lp_start[k]
        assert H.same(q.cells_dense(0), pairs[k].cells_dense(0))
        assert H.same(q.cells_dense(1), pairs[k].cells_dense(1))


def test_errors_are_reported_not_thrown(ctx):
    from historian_b200 import capi
    alphabet = "acgt"
    x = D.Profile.from_sequence("acgt", alphabet)
    bad = D.Profile(x.is_emit, x.trans_dest, x.trans_src, x.trans_lp, x.lp_absorb)  # src > dest
    with pytest.raises(capi.HistError) as e:
        ctx.upload(bad)
    assert e.value.code == -1 and "toposorted" in str(e.value)
    hmm20 = synth.random_hmm(np.random.default_rng(1), 20, 1)
    dx = ctx.upload(x)
    with pytest.raises(capi.HistError):
        ctx.prepare(dx, dx, hmm20)  # alphabet mismatch
    hmm = synth.random_hmm(np.random.default_rng(1), 4, 1)
    pair = ctx.prepare(dx, dx, hmm)
    with pytest.raises(capi.HistError) as e:
        pair.backward()  # BackwardMatrix needs a ForwardMatrix
    assert e.value.code == -4


def test_leaf_profiles_built_on_device_match_uploaded_ones(ctx):
    """hist_profile_from_sequences (Profile leaf constructor, profile.cpp:23-76) vs uploading the host-built
    flattened profile: identical DPMatrix tables, cells and likelihoods; wildcards, invalid characters,
    lower case, an empty sequence and ragged lengths."""
    rng = np.random.default_rng(5)
    alphabet = "arndcqeghilkmfpstwyv"
    seqs = ["".join(rng.choice(list(alphabet), size=L)) for L in (37, 5, 64, 1, 20)] + ["ar*dXc", "ARND", ""]
    dev = ctx.upload_sequences(seqs, alphabet, components=1)
    hmm = synth.random_hmm(rng, 20, 1, tl=0.2, tr=0.3)
    for a in range(0, len(seqs) - 1):
        b = a + 1
        hx, hy = D.Profile.from_sequence(seqs[a], alphabet, 1), D.Profile.from_sequence(seqs[b], alphabet, 1)
        ux, uy = ctx.upload(hx), ctx.upload(hy)
        p1, p2 = ctx.prepare(dev[a], dev[b], hmm), ctx.prepare(ux, uy, hmm)
        l1, l2 = p1.forward(), p2.forward()
        assert l1 == l2 or (np.isneginf(l1) and np.isneginf(l2))
        assert H.same(p1.cells_dense(0), p2.cells_dense(0))
        for t1, t2 in zip(p1.tables(), p2.tables()):
            assert H.same(t1, t2)


def test_queue_kernel_variant_still_matches():
    """HIST_B200_FILL=queue selects the in-CTA producer/consumer kernel (the fallback when the record buffer of
    the streamed path cannot be allocated); it must pass the same parity tests."""
    import subprocess
    import sys
    env = dict(os.environ, HIST_B200_FILL="queue")
    here = os.path.abspath(__file__)
    r = subprocess.run([sys.executable, "-m", "pytest", here, "-m", "gpu", "-q", "-x", "-k", "leaf_pair_banded or synthetic_dag or batch_fill"],
                       env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=os.path.dirname(os.path.dirname(here)))
    assert r.returncode == 0, r.stdout[-2000:]


def test_leaf_profiles_from_tokens_match_those_from_residues(ctx):
    """hist_profile_from_sequences (token strings) and hist_profile_from_residues (characters, tokenised inside the
    library as fastseq.cpp:10-16 does) build the same leaf profiles; hist_profiles_destroy / hist_pairs_destroy
    release a whole level."""
    rng = np.random.default_rng(9)
    alphabet = "arndcqeghilkmfpstwyv"
    seqs = ["".join(rng.choice(list(alphabet), size=L)) for L in (30, 7, 41)] + ["aRn*dZ"]
    toks = [np.array([alphabet.find(c.lower()) if c.lower() in alphabet else -1 for c in s], np.int8) for s in seqs]
    by_res = ctx.upload_sequences(seqs, alphabet, components=1)
    by_tok = ctx.upload_tokens(toks, len(alphabet), components=1)
    hmm = synth.random_hmm(rng, 20, 1, tl=0.3, tr=0.2)
    pairs = []
    for a, b in ((0, 1), (2, 3), (3, 0)):
        p1, p2 = ctx.prepare(by_res[a], by_res[b], hmm), ctx.prepare(by_tok[a], by_tok[b], hmm)
        assert p1.forward() == p2.forward()
        assert H.same(p1.cells_dense(0), p2.cells_dense(0))
        pairs += [p1, p2]
    ctx.close_all(pairs)
    ctx.close_all(by_res + by_tok)
    assert all(p.h is None for p in pairs + by_res + by_tok)


@pytest.mark.parametrize("case", ["lg_t1e-8", "ecmrest_t1e-6", "lg_far_apart"])
def test_posteriors_without_the_within_cell_range_exemption(ctx, case):
    """The five states of a cell share one exponent, so a state more than ~87 nats below its cell's maximum is stored as
    zero.  assert_log_close exempts such states; this test does NOT: on the inputs where the
    range is stretched furthest -- near-zero branch lengths (emission rows that are almost one-hot, transition weights
    like log(1e-10)), the 61-state codon model, and two unrelated sequences -- every (cell, state) whose exact posterior
    exp(F + B - lpEnd) is at least 1e-30 must be present with the right value, the posterior mass must be complete,
    and the best trace must be the exact-arithmetic one.  A flushed state could only matter if it carried posterior
    mass; this shows it never does."""
    from historian_b200 import model as M
    rng = np.random.default_rng(11)
    if case == "ecmrest_t1e-6":
        rm = M.RateModel.preset("ECMrest")
        x, y, (r1, r2) = synth.leaf_pair(rng, rm.alphabet, 40, 38, identity=0.8)
        hmm = M.pair_hmm(rm, 1e-6, 2e-6)
    elif case == "lg_t1e-8":
        rm = M.RateModel.preset("lg")
        x, y, (r1, r2) = synth.leaf_pair(rng, rm.alphabet, 90, 85, identity=0.9)
        hmm = M.pair_hmm(rm, 1e-8, 1e-8)
    else:
        rm = M.RateModel.preset("lg")
        x, y, (r1, r2) = synth.leaf_pair(rng, rm.alphabet, 70, 75, identity=0.05)
        hmm = M.pair_hmm(rm, 0.02, 0.03)
    env = D.Envelope.from_guide_rows(r1, r2, 6)
    pair, dx, dy = prepare(ctx, x, y, hmm, env)
    lp_end = pair.forward()
    pair.backward()
    gf, gb = pair.cells_dense(0), pair.cells_dense(1)
    O.set_exact_lse(True)
    try:
        ef, elp = O.forward(x, y, hmm, env)
        eb, _ = O.backward(x, y, hmm, env)
        ebest = O.trace_best(x, y, hmm, env, np.ascontiguousarray(ef), (x.n_states - 1, y.n_states - 1, D.EEE))
    finally:
        O.set_exact_lse(False)
    assert np.isfinite(elp) and abs(lp_end - elp) <= REL_EXACT * abs(elp) + ABS_EXACT
    with np.errstate(invalid="ignore"):
        epost, gpost = ef + eb - elp, gf + gb - lp_end
    live = np.isfinite(epost) & (epost >= np.log(1e-30))
    assert live.sum() > 100
    assert np.all(np.isfinite(gpost[live])), "%d (cell, state) entries with posterior >= 1e-30 were flushed" % (~np.isfinite(gpost[live])).sum()
    assert np.max(np.abs(gpost[live] - epost[live])) <= 5e-4, np.max(np.abs(gpost[live] - epost[live]))
    # every anti-diagonal... the posterior mass of the cells of one x row that absorb or delete it sums to one
    big = np.isfinite(epost) & (epost >= np.log(1e-9))
    assert np.max(np.abs(np.exp(gpost[big]) - np.exp(epost[big]))) <= 1e-4
    if case != "lg_t1e-8":
        # (at t = 1e-8 the two orders of an adjacent insertion and deletion have the same weight to ~1e-8: the arg-max
        # between them is decided below FP32 resolution, north_star's "wherever the margin exceeds the tolerance")
        assert H.same(pair.trace_best(), ebest)


def test_handle_batches_equal_the_object_api(ctx):
    """upload_packed / PrepareBatch.run_batch / fill_batch on HandleBatch (no Python object per profile or pair) give the same
    likelihoods as the per-object calls: the same C-ABI entry points underneath."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    from historian_b200 import capi
    lg, guide = bench.make_family(3, n_leaves=16)
    hmm, inputs = bench.cherry_inputs(lg, guide, 8)
    seqs = bench.cherry_sequences(guide, 8)
    leaves = ctx.upload_sequences(seqs, lg.alphabet, lg.components)
    plan = capi.PrepareBatch(ctx, leaves[0::2], leaves[1::2], hmm, [e for _, _, e in inputs], backward=True)
    want = ctx.fill_batch(plan.run(), backward=True)
    packed = np.frombuffer(b"".join(seqs), np.uint8).copy()
    off = np.concatenate([[0], np.cumsum([len(q) for q in seqs])]).astype(np.int64)
    for _ in range(2):
        lb = ctx.upload_packed(packed, off, lg.alphabet, lg.components)
        pb = plan.run_batch(lb)
        got = ctx.fill_batch(pb, backward=True)
        assert len(lb) == 16 and len(pb) == 8
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
        pb.close()
        lb.close()


def test_two_contexts_with_fills_in_flight(ctx):
    """hist_fill_batch with lp_end == NULL + hist_fill_batch_results, alternating between two contexts on the device (the
    pipelined end-to-end step of bench.py): same likelihoods as the synchronous call; results of unfilled pairs are refused."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    from historian_b200 import capi
    lg, guide = bench.make_family(3, n_leaves=16)
    hmm, inputs = bench.cherry_inputs(lg, guide, 8)
    seqs = bench.cherry_sequences(guide, 8)
    packed = np.frombuffer(b"".join(seqs), np.uint8).copy()
    off = np.concatenate([[0], np.cumsum([len(q) for q in seqs])]).astype(np.int64)
    envs = [e for _, _, e in inputs]
    other = capi.Context(0)
    try:
        lanes = []
        for c in (ctx, other):
            leaves = c.upload_sequences(seqs, lg.alphabet, lg.components)
            lanes.append((c, capi.PrepareBatch(c, leaves[0::2], leaves[1::2], hmm, envs, backward=True), leaves))
        want = ctx.fill_batch(lanes[0][1].run(), backward=True)
        jobs = []
        for k in range(4):
            c, plan, _ = lanes[k % 2]
            lb = c.upload_packed(packed, off, lg.alphabet, lg.components)
            pb = plan.run_batch(lb)
            if k == 0:
                with pytest.raises(capi.HistError) as e:
                    c.fill_results(pb, backward=True)
                assert e.value.code == -4
            c.fill_batch_async(pb, backward=True)
            jobs.append((c, lb, pb))
            if len(jobs) == 2:
                c0, lb0, pb0 = jobs.pop(0)
                got = c0.fill_results(pb0, backward=True)
                assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
                pb0.close()
                lb0.close()
        c0, lb0, pb0 = jobs.pop(0)
        got = c0.fill_results(pb0, backward=True)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
        pb0.close()
        lb0.close()
        for _, plan, leaves in lanes:
            del plan
            for p in leaves:
                p.close()
    finally:
        other.close()
