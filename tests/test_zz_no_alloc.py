"""hist_profile_sum_path_absorb_noalloc: the allocation-free variant of P3 (Profile::calcSumPathAbsorbProbs,
profile.cpp:112-131) against the existing entry point hist_profile_sum_path_absorb, through the C ABI.

The entry point has no counterpart in the reference (it was added for callers
that cannot call cudaMalloc).  This file is named so that pytest collects it last."""
import numpy as np
import pytest

import synth
from historian_b200 import capi

pytestmark = pytest.mark.gpu


def _compare(ctx, p, logw, logins):
    dp = ctx.upload(p)
    try:
        want, want_cum = dp.sum_path_absorb(logw, logins)
        got, cum = dp.sum_path_absorb(logw, logins, no_alloc=True)
    finally:
        dp.close()
    # same terms, same fold order per state (in-lists ascending in transition index, lane-strided): the tolerance of
    # tests/test_elim_gpu.py::test_sum_path_absorb_on_device
    assert np.isfinite(want)
    assert abs(got - want) <= 1e-9 * abs(want) + 1e-9
    fin = np.isfinite(want_cum)
    assert np.array_equal(np.isfinite(cum), fin)
    assert np.all(np.abs(cum[fin] - want_cum[fin]) <= 1e-9 * np.abs(want_cum[fin]) + 1e-9)
    return got


def test_noalloc_matches_existing_entry_point_on_two_calls_in_a_row(ctx):
    """Two calls in a row on different profiles -- the second one smaller, with another alphabet and component count,
    so that anything the first call left in the shared scratch array would show -- and then the first again."""
    rng = np.random.default_rng(41)
    cases = []
    for A, Cc, n in ((20, 2, 400), (4, 1, 37), (61, 1, 90)):
        p = synth.random_profile(rng, n, A, Cc)
        cases.append((p, np.log(rng.dirichlet(np.ones(Cc))), np.log(rng.dirichlet(np.ones(A), size=Cc))))
    first = _compare(ctx, *cases[0])
    _compare(ctx, *cases[1])
    _compare(ctx, *cases[2])
    again = _compare(ctx, *cases[0])
    assert again == first   # one block, fixed fold order: bit-reproducible


def test_noalloc_on_leaf_chains(ctx):
    rng = np.random.default_rng(42)
    for L in (1, 50, 3000):
        seq = "".join(rng.choice(list("acgt"), L))
        (dx,) = ctx.upload_sequences([seq], "acgt")
        logw, logins = np.zeros(1), np.log(np.full((1, 4), 0.25))
        want, _ = dx.sum_path_absorb(logw, logins)
        got, _ = dx.sum_path_absorb(logw, logins, no_alloc=True)
        dx.close()
        assert abs(got - want) <= 1e-9 * abs(want)
        assert abs(got - L * np.log(0.25)) <= 1e-9 * abs(want)   # every residue contributes log(1/4)


def test_noalloc_rejects_profiles_beyond_the_scratch_array(ctx):
    limit = ctx.lib.hist_noalloc_scratch_bytes()
    assert limit == 4 << 20
    L = limit // 40   # ~45 B of scratch per state of a chain: well beyond the array
    (dx,) = ctx.upload_sequences(["a" * L], "acgt")
    logw, logins = np.zeros(1), np.log(np.full((1, 4), 0.25))
    with pytest.raises(capi.HistError) as e:
        dx.sum_path_absorb(logw, logins, no_alloc=True)
    assert e.value.code == -5 and "scratch" in str(e.value)
    want, _ = dx.sum_path_absorb(logw, logins)   # the allocating entry point still serves it
    assert abs(want - L * np.log(0.25)) <= 1e-9 * abs(want)
    # and the shared array is usable afterwards
    (dy,) = ctx.upload_sequences(["acgtacgt"], "acgt")
    got, _ = dy.sum_path_absorb(logw, logins, no_alloc=True)
    assert abs(got - 8 * np.log(0.25)) <= 1e-9 * abs(got)
    dx.close()
    dy.close()
