"""GPU parity tests of hist_profile_eliminate (numeric core of ForwardMatrix::makeProfile,
forward.cpp:686-843) against the oracle, on the matrix the GPU filled.

State selection, state numbering, transition structure and best-path via lists are integer
bookkeeping: bit-exact.  lpPath / lpBestAlignPath / absorb tables / fwdLogProb are doubles built from
the same inputs by exact log-sum-exp on both sides: |diff| <= 1e-9."""
import numpy as np
import pytest

import helpers as H
import oracle_lib as O
import synth
from historian_b200 import descs as D
from test_trace_gpu import filled_pair, mt_seed

pytestmark = pytest.mark.gpu

CASES = H.golden_cases()
TOL = 1e-9


def cell_set(*paths):
    s = set()
    for p in paths:
        s.update(tuple(int(v) for v in c) for c in p)
    return np.array(sorted(s), np.int32).reshape(-1, 3)


def compare(g, o, what, tol=TOL):
    for k in ("state_cell", "trans_src", "trans_dest", "via_off", "via"):
        assert H.same(g[k], o[k]), "%s: %s differs (%s vs %s)" % (what, k, g[k].shape, o[k].shape)
    for k in ("state_lp_absorb", "state_fwd_lp", "trans_lp", "trans_best_lp"):
        a, b = np.asarray(g[k]), np.asarray(o[k])
        assert H.same(np.isneginf(a), np.isneginf(b)), (what, k)
        fin = np.isfinite(b)
        assert np.all(np.abs(a[fin] - b[fin]) <= tol * (1 + np.abs(b[fin]))), (what, k, np.abs(a[fin] - b[fin]).max())


def check_elim(ctx, x, y, hmm, env, what, seed=5489):
    pair, keep, lp_end = filled_pair(ctx, x, y, hmm, env)
    if not np.isfinite(lp_end):
        pytest.skip("zero likelihood")
    fwd = pair.cells_dense(0)
    best = pair.trace_best()
    samples = pair.trace_sample(mt_seed(seed), 5)
    post, _ = pair.posterior_cells(0.05)
    end = np.array([[x.n_states - 1, y.n_states - 1, D.EEE]], np.int32)
    sets = {"best": cell_set(best), "best+samples": cell_set(best, *samples), "posterior": cell_set(best, post, end)}
    O.set_exact_lse(True)
    try:
        for name, cells in sets.items():
            for strategy in (D.COLLAPSE_CHAINS, 0, D.COLLAPSE_CHAINS | D.KEEP_GAPS_OPEN):
                o = O.profile_eliminate(x, y, hmm, env, fwd, lp_end, cells, strategy)
                g = pair.eliminate(cells, strategy)
                # one-component codon alphabets: the leftMultiply tables come from FP32 tensor arithmetic (hist_codon.cu)
                tensor = x.components == 1 and 32 < x.alph_size <= 64
                compare(g, o, "%s/%s/strategy%d" % (what, name, strategy), 2e-6 if tensor else TOL)
    finally:
        O.set_exact_lse(False)
    return pair


@pytest.mark.parametrize("cid,path,node", CASES[1::3], ids=[c[0] for c in CASES[1::3]])
def test_eliminate_on_golden_inputs(ctx, cid, path, node):
    x, y, hmm, env = H.node_inputs(H.load_golden(path), node)
    check_elim(ctx, x, y, hmm, env, cid)


@pytest.mark.parametrize("seed", range(6))
def test_eliminate_on_synthetic_dags(ctx, seed):
    rng = np.random.default_rng(3000 + seed)
    A = [4, 20, 4, 20, 7, 4][seed]
    C = [1, 1, 2, 1, 3, 1][seed]
    x = synth.random_profile(rng, int(rng.integers(8, 50)), A, C, one_hot=seed % 2 == 0, jitter_leaf=seed >= 3)
    y = synth.random_profile(rng, int(rng.integers(8, 50)), A, C, one_hot=seed % 2 == 0, jitter_leaf=seed >= 3)
    hmm = synth.random_hmm(rng, A, C)
    env = D.Envelope(-1) if seed % 3 == 0 else synth.random_envelope(rng, x, y, int(rng.integers(2, 8)))
    check_elim(ctx, x, y, hmm, env, "synth%d" % seed, seed=7 + seed)


def test_eliminate_leaf_pair_300(ctx):
    rng = np.random.default_rng(23)
    x, y, (r1, r2) = synth.leaf_pair(rng, "arndcqeghilkmfpstwyv", 300, 290)
    hmm = synth.random_hmm(rng, 20, 1, tl=0.1, tr=0.1)
    check_elim(ctx, x, y, hmm, D.Envelope.from_guide_rows(r1, r2, 20), "leaf300")


def test_eliminate_argument_checks(ctx):
    from historian_b200 import capi
    x, y = D.Profile.from_sequence("acgtac", "acgt"), D.Profile.from_sequence("acgac", "acgt")
    hmm = synth.random_hmm(np.random.default_rng(1), 4, 1)
    pair, keep, _ = filled_pair(ctx, x, y, hmm, D.Envelope(-1), backward=False)
    best = pair.trace_best()
    with pytest.raises(capi.HistError) as e:
        pair.eliminate(best[::-1])  # not ascending
    assert e.value.code == -1
    with pytest.raises(capi.HistError):
        pair.eliminate(best[:-1])  # end cell missing


def test_sum_path_absorb_on_device(ctx):
    """P3, Profile::calcSumPathAbsorbProbs (profile.cpp:112-131) through hist_profile_sum_path_absorb vs the oracle
    with exact log-sum-exp (1e-9 relative; the device and the oracle fold the same terms, the device in a
    lane-strided order), on leaf chains and on random DAG profiles with several components."""
    import oracle_lib as O
    import synth
    rng = np.random.default_rng(12)
    O.set_exact_lse(True)
    try:
        for A, Cc, n in ((20, 1, 60), (4, 2, 35), (20, 3, 150), (61, 1, 40)):
            p = synth.random_profile(rng, n, A, Cc)
            logw = np.log(rng.dirichlet(np.ones(Cc)))
            logins = np.log(rng.dirichlet(np.ones(A), size=Cc))
            want, want_cum = O.sum_path_absorb(p, logw, logins)
            dp = ctx.upload(p)
            got, cum = dp.sum_path_absorb(logw, logins)
            assert abs(got - want) <= 1e-9 * abs(want) + 1e-9
            fin = np.isfinite(want_cum)
            assert np.array_equal(np.isfinite(cum), fin)
            assert np.all(np.abs(cum[fin] - want_cum[fin]) <= 1e-9 * np.abs(want_cum[fin]) + 1e-9)
            dp.close()
        x, _, _ = synth.leaf_pair(rng, "acgt", 50, 40)
        logw, logins = np.zeros(1), np.log(np.full((1, 4), 0.25))
        want, _ = O.sum_path_absorb(x, logw, logins)
        dx = ctx.upload(x)
        got, _ = dx.sum_path_absorb(logw, logins)
        assert abs(got - want) <= 1e-9 * abs(want)
        assert abs(got - 50 * np.log(0.25)) <= 1e-9 * abs(want)   # every residue contributes log(1/4)
        dx.close()
    finally:
        O.set_exact_lse(False)
