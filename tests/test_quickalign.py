"""Guide-alignment stage: QuickAlignMatrix (quickalign.cpp:7-214) inside a DiagonalEnvelope.

Fixtures tests/golden/quickalign_*.npz come from the REAL reference (oracle/ref_dump_quickalign.cpp through
tools/make_golden_quickalign.py): full envelopes (the reference's own testquickalign input, data/PF16593.pair.fa, and a
simulated LG family) and sparse k-mer-seeded envelopes.  Everything is compared for EQUALITY: Viterbi has no
log-sum-exp, the device adds the same doubles in the same order (unfused), so scores are bit-equal and paths identical."""
import glob
import os

import numpy as np
import pytest

import helpers as H
import oracle_lib as O

FIX = sorted(glob.glob(os.path.join(H.GOLDEN_DIR, "quickalign_*.npz")))
IDS = [os.path.basename(f)[11:-4] for f in FIX]


def cases(d):
    A = int(round(np.sqrt(len(d["submat"]))))
    sub = d["submat"].reshape(A, A)
    for k in range(int(d["n_pairs"][0])):
        p = "p%d." % k
        yield k, (d[p + "x_tok"], d[p + "y_tok"], d[p + "diagonals"]), sub, d["scores"], float(d[p + "result"][0]), tuple(int(v) for v in d[p + "end"]), d.get(p + "path")


@pytest.mark.parametrize("path", FIX, ids=IDS)
def test_oracle_restatement_equals_the_reference(path):
    d = H.load_golden(path)
    for k, pair, sub, scores, want, end, wpath in cases(d):
        got, gend, gpath = O.quickalign(*pair, sub, scores)
        assert got == want and gend == end, (k, got, want, gend, end)
        assert np.array_equal(gpath, wpath), k


def test_reference_golden_alignment_of_pf16593_pair():
    """The reference's own `make testquickalign` golden (data/testquickalign.out.fa) is the gapped form of this path."""
    d = H.load_golden(os.path.join(H.GOLDEN_DIR, "quickalign_pf16593_pair_full.npz"))
    ops = d["p0.path"]
    assert int(np.sum(ops & 1 > 0)) == len(d["p0.x_tok"]) and int(np.sum(ops & 2 > 0)) == len(d["p0.y_tok"])
    # testquickalign.out.fa: 13 aligned columns, a 3-column gap in the first row, then aligned to the end
    assert "".join(map(str, ops)) == "3" * 13 + "2" * 3 + "3" * 20


@pytest.mark.gpu
@pytest.mark.parametrize("path", FIX, ids=IDS)
def test_device_equals_the_reference(ctx, path):
    d = H.load_golden(path)
    cs = list(cases(d))
    res, ends, paths = ctx.quickalign([c[1] for c in cs], cs[0][2], cs[0][3])
    for (k, _, _, _, want, end, wpath), got, gend, gpath in zip(cs, res, ends, paths):
        assert got == want, (k, got, want)
        assert tuple(gend) == end, (k, gend, end)
        assert np.array_equal(gpath, wpath), k


def random_case(rng, A, lx, ly, band, indel=0.04):
    """Two related sequences and a banded envelope around a few seed diagonals; score table of the LG fixture's shape."""
    x = rng.integers(0, A, lx).astype(np.int8)
    y = []
    for t in x:
        r = rng.random()
        if r < indel:
            continue
        y.append(t if rng.random() < 0.7 else rng.integers(0, A))
        if rng.random() < indel:
            y.extend(rng.integers(0, A, rng.integers(1, 4)))
    y = np.array(y[:ly] if len(y) >= 2 else [0, 1], np.int8)
    x[rng.random(lx) < 0.01] = -1                      # wildcards
    lo, hi = 1 - len(y), lx - 1
    seeds = [0, int(lx - len(y))] + [int(v) for v in rng.integers(lo, hi + 1, 2)]
    dg = sorted({v for s in seeds for v in range(max(lo, s - band), min(hi, s + band) + 1)})
    return x, y, np.array(dg, np.int32)


@pytest.mark.gpu
def test_device_equals_the_oracle_on_random_envelopes(ctx):
    """Ragged batch: lengths 2 .. 1000, bands 0 .. 64, gapped diagonal sets (several disjoint bands), wildcards; and a
    full 300 x 280 envelope.  Scores bit-equal, end cells and paths identical."""
    d = H.load_golden(os.path.join(H.GOLDEN_DIR, "quickalign_lg6x120_full.npz"))
    sub, scores = d["submat"].reshape(20, 20), d["scores"]
    rng = np.random.default_rng(5)
    pairs = [random_case(rng, 20, int(lx), int(lx), int(b)) for lx, b in ((2, 0), (5, 1), (40, 3), (200, 8), (333, 20), (1000, 32), (1000, 64), (700, 0))]
    x, y, _ = random_case(rng, 20, 300, 280, 4)
    pairs.append((x, y, np.arange(1 - len(y), len(x), dtype=np.int32)))
    res, ends, paths = ctx.quickalign(pairs, sub, scores)
    n_finite = 0
    for k, pr in enumerate(pairs):
        want, end, wpath = O.quickalign(*pr, sub, scores)
        assert res[k] == want or (np.isinf(want) and np.isinf(res[k])), (k, res[k], want)
        assert tuple(ends[k]) == end, (k, ends[k], end)
        if np.isfinite(want):
            n_finite += 1
            assert np.array_equal(paths[k], wpath), k
            assert int(np.sum(paths[k] & 1 > 0)) == len(pr[0]) and int(np.sum(paths[k] & 2 > 0)) == len(pr[1])
        else:
            assert len(paths[k]) == 0
    assert n_finite >= 7


@pytest.mark.gpu
def test_argument_checks(ctx):
    from historian_b200 import capi
    d = H.load_golden(os.path.join(H.GOLDEN_DIR, "quickalign_pf16593_pair_full.npz"))
    sub, scores = d["submat"].reshape(20, 20), d["scores"]
    x, y = d["p0.x_tok"], d["p0.y_tok"]
    with pytest.raises(capi.HistError):
        ctx.quickalign([(x, y, np.array([3, 2], np.int32))], sub, scores)          # not ascending
    with pytest.raises(capi.HistError):
        ctx.quickalign([(x, y, np.array([len(x)], np.int32))], sub, scores)        # outside [1 - yLen, xLen - 1]


def test_host_envelope_and_scores_match_the_reference():
    """historian_b200/guide.py: the k-mer seeded envelope (diagenv.cpp:101-190, fixed threshold) reproduces the reference's
    diagonal lists exactly, and the constructor's score table (quickalign.cpp:22-50) its values to rounding."""
    from historian_b200 import guide
    from historian_b200.model import RateModel
    d = H.load_golden(os.path.join(H.GOLDEN_DIR, "quickalign_pf16593_pair_full.npz"))
    sub, sc = guide.quickalign_scores(RateModel.preset("lg"), 1.0)
    assert np.allclose(sub.ravel(), d["submat"], rtol=0, atol=1e-12) and np.allclose(sc, d["scores"], rtol=0, atol=1e-12)
    for name, A, args in (("quickalign_lg8x300_sparse", 20, (5, 16, 2)), ("quickalign_jc5x200_sparse", 4, (6, 8, 1))):
        d = H.load_golden(os.path.join(H.GOLDEN_DIR, name + ".npz"))
        for k in range(int(d["n_pairs"][0])):
            p = "p%d." % k
            dg = guide.sparse_envelope(d[p + "x_tok"], d[p + "y_tok"], A, *args)
            assert np.array_equal(dg, d[p + "diagonals"]), (name, k)
            assert guide.envelope_cells(dg, len(d[p + "x_tok"]), len(d[p + "y_tok"])) > 0


@pytest.mark.gpu
def test_full_envelope_without_the_diagonal_list(ctx):
    """diagonals == NULL stands for initFull: same scores, end cells and paths as the explicit list."""
    d = H.load_golden(os.path.join(H.GOLDEN_DIR, "quickalign_lg6x120_full.npz"))
    cs = list(cases(d))
    res, ends, paths = ctx.quickalign([(c[1][0], c[1][1], None) for c in cs], cs[0][2], cs[0][3])
    for (k, _, _, _, want, end, wpath), got, gend, gpath in zip(cs, res, ends, paths):
        assert got == want and tuple(gend) == end and np.array_equal(gpath, wpath), k
    with pytest.raises(Exception):   # a NULL list with another count is rejected
        from historian_b200 import descs as D
        import ctypes as C
        x, y = cs[0][1][0].astype(np.int8), cs[0][1][1].astype(np.int8)
        sub = np.ascontiguousarray(cs[0][2], np.float64)
        sc = D.QuickAlignScores(int(sub.shape[0]), sub.ctypes.data_as(C.POINTER(C.c_double)), *[float(v) for v in cs[0][3]])
        pr = (D.QuickAlignPair * 1)(D.QuickAlignPair(len(x), len(y), x.ctypes.data_as(C.POINTER(C.c_int8)), y.ctypes.data_as(C.POINTER(C.c_int8)), 5, None))
        r = np.zeros(1)
        ctx.check(ctx.lib.hist_quickalign_batch(ctx.h, 1, pr, C.byref(sc), r.ctypes.data_as(C.POINTER(C.c_double)), None, None, None, None, None))


@pytest.mark.parametrize("name,band,klen,thr,limit", [("quickalign_lg8x300_sparse", 16, 5, 2, 1 << 40), ("quickalign_jc5x200_sparse", 8, 6, 1, 1 << 40),
                                                      ("quickalign_lg6x120_full", 64, 6, -1, 1 << 40), ("quickalign_lg6x250_kmatchmb", 12, 5, -1, 400000)])
def test_library_envelopes_equal_the_reference(name, band, klen, thr, limit):
    """DiagonalEnvelope::initSparse / initFull in the library's host code (hist_diagonal_envelope, what hist_guide_align builds per
    pair) against the diagonals the reference built for the fixtures' pairs; and against the Python host restatement
    (guide.sparse_envelope).  No device work."""
    from historian_b200 import capi, guide
    d = H.load_golden(os.path.join(H.GOLDEN_DIR, name + ".npz"))
    A = int(round(np.sqrt(len(d["submat"]))))
    for k, (x, y, want), *_ in cases(d):
        # threshold -1 with a memory limit above xLen * yLen * cellSize: the full envelope, the reference's default
        # (kmatchmb: -1 with a limit BELOW it: the largest set of seeded bands whose storage diagonals fit, diagenv.cpp:158-186)
        got = capi.diagonal_envelope(x, y, A, sparse=True, kmer_len=klen, kmer_threshold=thr, band_size=band, cell_size=40, max_size=limit)
        assert np.array_equal(got, want), k
        if thr >= 0:
            assert np.array_equal(guide.sparse_envelope(x, y, A, kmer_len=klen, band=band, threshold=thr), want), k
    # memory-bounded threshold (-kmatchmb): shrinking the limit can only drop seeded diagonals, diagonal 0 always stays
    _, (x, y, _), *_ = next(iter(cases(d)))
    sizes = [len(capi.diagonal_envelope(x, y, A, sparse=True, kmer_len=5, kmer_threshold=-1, band_size=8, cell_size=40, max_size=m))
             for m in (0, 20000, 200000, 2000000, 1 << 40)]
    assert sizes[0] == 1 and sizes == sorted(sizes) and sizes[-1] == len(x) + len(y) - 1
    assert 0 in capi.diagonal_envelope(x, y, A, sparse=True, kmer_len=5, kmer_threshold=-1, band_size=8, cell_size=40, max_size=20000)
