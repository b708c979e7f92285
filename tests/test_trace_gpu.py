"""GPU parity tests of the traceback kernels (hist_trace_best / hist_trace_sample /
hist_trace_best_forward) through the C ABI.

The oracle (bestTrace / sampleTrace / BackwardMatrix::bestTrace restated from forward.cpp:225-302,
1287-1300) walks the SAME matrix the GPU filled (the dense natural-log cells from hist_cells_fetch),
with the reference's exact log-sum-exp for the emission terms.  Paths are integer data: the bar is
bit-exact equality, including the generator state after sampling (std::mt19937 draws through
libstdc++'s uniform_real_distribution, two 32-bit outputs per step)."""
import numpy as np
import pytest

import helpers as H
import oracle_lib as O
import synth
from historian_b200 import descs as D

pytestmark = pytest.mark.gpu

CASES = H.golden_cases()


def mt_seed(seed=5489):
    """std::mt19937(seed) state image: 624 words + position (624 = 'twist before the first draw')."""
    mt = np.zeros(625, np.uint32)
    x = np.uint64(seed & 0xffffffff)
    mt[0] = x
    for k in range(1, 624):
        x = (np.uint64(1812433253) * (x ^ (x >> np.uint64(30))) + np.uint64(k)) & np.uint64(0xffffffff)
        mt[k] = x
    mt[624] = 624
    return mt


def filled_pair(ctx, x, y, hmm, env, backward=True):
    dx, dy = ctx.upload(x), ctx.upload(y)
    pair = ctx.prepare(dx, dy, hmm, env)
    lp_end = pair.forward()
    if backward:
        pair.backward()
    return pair, (dx, dy), lp_end


def check_traces(ctx, x, y, hmm, env, what, n_samples=3, seed=5489):
    pair, keep, lp_end = filled_pair(ctx, x, y, hmm, env)
    if not np.isfinite(lp_end):
        pytest.skip("zero likelihood")
    fwd, bwd = pair.cells_dense(0), pair.cells_dense(1)
    end = (x.n_states - 1, y.n_states - 1, D.EEE)
    O.set_exact_lse(True)
    try:
        obest = O.trace_best(x, y, hmm, env, fwd, end)
        gbest = pair.trace_best()
        assert H.same(gbest, obest), "%s: best trace differs at step %s" % (what, first_diff(gbest, obest))
        # traces from interior cells, both directions (postProbProfile's addTrace, forward.cpp:1356-1379)
        rng = np.random.default_rng(1)
        for k in rng.choice(len(obest) - 1, size=min(4, len(obest) - 1), replace=False):
            c = tuple(int(v) for v in obest[k])
            assert H.same(pair.trace_best(c), O.trace_best(x, y, hmm, env, fwd, c)), (what, "back from", c)
            assert H.same(pair.trace_best_forward(c), O.trace_best_forward(x, y, hmm, env, bwd, c)), (what, "forward from", c)
        # posterior decoding (cellsAbovePostProbThreshold, forward.cpp:1302-1319): same doubles -> same order
        for thr in (0.5, 0.01):
            oc, olpp = O.posterior_cells(x, y, hmm, env, fwd, bwd, lp_end, thr)
            gc, glpp = pair.posterior_cells(thr)
            assert H.same(gc, oc), "%s: posterior cell list differs (threshold %g): %d vs %d" % (what, thr, len(gc), len(oc))
            assert H.same(glpp, olpp)
        st_o, st_g = mt_seed(seed), mt_seed(seed)
        otr, st_o = O.trace_sample(x, y, hmm, env, fwd, st_o, n_samples)
        gtr = pair.trace_sample(st_g, n_samples)
        for k in range(n_samples):
            assert H.same(gtr[k], otr[k]), "%s: sampled trace %d differs at step %s" % (what, k, first_diff(gtr[k], otr[k]))
        assert H.same(st_g, st_o), what + ": generator state after sampling differs"
    finally:
        O.set_exact_lse(False)
    return pair


def first_diff(a, b):
    n = min(len(a), len(b))
    for k in range(n):
        if not np.array_equal(a[k], b[k]):
            return k, a[k].tolist(), b[k].tolist()
    return n, len(a), len(b)


def test_mt19937_image_matches_numpy():
    """The seeded state image equals numpy's MT19937 (same init_genrand) - guards the test helper."""
    st = np.random.MT19937()
    st._legacy_seeding(5489)
    assert np.array_equal(st.state["state"]["key"], mt_seed(5489)[:624])


@pytest.mark.parametrize("cid,path,node", CASES[::3], ids=[c[0] for c in CASES[::3]])
def test_traces_on_golden_inputs(ctx, cid, path, node):
    x, y, hmm, env = H.node_inputs(H.load_golden(path), node)
    check_traces(ctx, x, y, hmm, env, cid)


@pytest.mark.parametrize("seed", range(6))
def test_traces_on_synthetic_dags(ctx, seed):
    rng = np.random.default_rng(2000 + seed)
    A = [4, 20, 4, 20, 7, 4][seed]
    C = [1, 1, 2, 1, 3, 1][seed]
    x = synth.random_profile(rng, int(rng.integers(8, 60)), A, C, one_hot=seed % 2 == 0, jitter_leaf=seed >= 3)
    y = synth.random_profile(rng, int(rng.integers(8, 60)), A, C, one_hot=seed % 2 == 0, jitter_leaf=seed >= 3)
    hmm = synth.random_hmm(rng, A, C)
    env = D.Envelope(-1) if seed % 3 == 0 else synth.random_envelope(rng, x, y, int(rng.integers(2, 8)))
    check_traces(ctx, x, y, hmm, env, "synth%d" % seed, n_samples=4, seed=1 + seed)


def test_traces_leaf_pair_banded_600(ctx):
    """Enough steps (~1200 per trace x 10 traces) to cross several generator twists (624 words)."""
    rng = np.random.default_rng(17)
    x, y, (r1, r2) = synth.leaf_pair(rng, "arndcqeghilkmfpstwyv", 600, 590)
    hmm = synth.random_hmm(rng, 20, 1, tl=0.05, tr=0.05)
    env = D.Envelope.from_guide_rows(r1, r2, 20)
    check_traces(ctx, x, y, hmm, env, "leaf600", n_samples=10)


def test_trace_requires_fill_and_capacity(ctx):
    from historian_b200 import capi
    import ctypes as C
    x, y = D.Profile.from_sequence("acgtac", "acgt"), D.Profile.from_sequence("acgac", "acgt")
    hmm = synth.random_hmm(np.random.default_rng(1), 4, 1)
    dx, dy = ctx.upload(x), ctx.upload(y)
    pair = ctx.prepare(dx, dy, hmm)
    with pytest.raises(capi.HistError) as e:
        pair.trace_best()
    assert e.value.code == -4  # HIST_ERR_STATE
    pair.forward()
    buf = (D.Cell * 2)()
    n = C.c_int64()
    rc = ctx.lib.hist_trace_best(pair.h, D.Cell(6 + 1, 5 + 1, D.EEE), buf, 2, C.byref(n))
    assert rc == -5  # HIST_ERR_CAPACITY
