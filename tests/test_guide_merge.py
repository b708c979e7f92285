"""Guide-alignment stage as a whole: AlignGraph (span.cpp:35-158) = trial edges, QuickAlignMatrix per
edge, spanning tree, alignPathMerge (alignpath.cpp:98-207).

Fixtures tests/golden/guide_*.npz come from the REAL reference (oracle/ref_dump_guide.cpp through
tools/make_golden_guide.py): the random graph with the default and another seed, the all-pairs graph, full envelopes, a
fixed k-mer threshold (-kmatchn) and the memory-bounded threshold (-kmatchmb).  Everything is compared for EQUALITY:
the guide alignment, the edges in the order they were drawn and their scores.

CPU tests: the library's host-only merge (hist_align_path_merge) against oracle/align_path_merge.py (the reference's loops,
restated) on the fixtures' own spanning-tree alignments and on random families of overlapping alignments."""
import glob
import os
import sys

import numpy as np
import pytest

import helpers as H
import oracle_lib as O
from historian_b200 import capi

sys.path.insert(0, O.ORACLE_DIR)
from align_path_merge import align_path_merge as oracle_merge   # noqa: E402

FIX = sorted(glob.glob(os.path.join(H.GOLDEN_DIR, "guide_*.npz")))
IDS = [os.path.basename(f)[6:-4] for f in FIX]


def _text(a):
    return bytes(a).decode()


def _bits(gapped):
    return np.array([[c != "-" for c in row] for row in gapped], np.uint8)


def _project(bits, rows):
    """The sub-alignment of some rows of an alignment, empty columns removed, as (rows, bits)."""
    sub = bits[rows]
    return list(rows), sub[:, sub.any(axis=0)]


def _as_paths(alignments):
    return [{int(r): [bool(v) for v in b[k]] for k, r in enumerate(rows)} for rows, b in alignments]


def _oracle_bits(alignments, n_rows):
    out = oracle_merge(_as_paths(alignments))
    cols = len(next(iter(out.values()))) if out else 0
    m = np.zeros((n_rows, cols), np.uint8)
    for r, p in out.items():
        m[r] = p
    return m


@pytest.mark.parametrize("path", FIX, ids=IDS)
def test_merge_of_projected_pairs_equals_the_oracle_and_keeps_every_pair(path):
    """Pairwise projections of the reference's guide alignment along a random spanning tree: the library's merge equals the
    restated reference loops, and every input alignment is a projection of the result."""
    d = H.load_golden(path)
    bits = _bits(_text(d["gapped"]).split("\n"))
    n = len(bits)
    rng = np.random.default_rng(n)
    order = rng.permutation(n)
    als = [_project(bits, sorted((int(order[k]), int(order[rng.integers(0, k)])))) for k in range(1, n)]
    lens = bits.sum(axis=1)
    got = capi.align_path_merge(als, lens)
    assert np.array_equal(got, _oracle_bits(als, n))
    assert np.array_equal(got.sum(axis=1), lens)
    for rows, b in als:
        assert np.array_equal(_project(got, rows)[1], b)


def test_merge_of_overlapping_multi_row_alignments():
    rng = np.random.default_rng(5)
    for trial in range(30):
        n, cols = int(rng.integers(2, 9)), int(rng.integers(1, 40))
        truth = (rng.random((n, cols)) < 0.6).astype(np.uint8)
        truth[0, truth.sum(axis=0) == 0] = 1
        truth[:, 0] |= truth.sum(axis=1) == 0     # no empty rows
        als, covered = [], {0}
        while len(covered) < n:   # connected family of 2..4-row alignments
            rows = {int(rng.choice(sorted(covered)))} | {int(v) for v in rng.integers(0, n, int(rng.integers(1, 4)))}
            if len(rows) < 2:
                continue
            covered |= rows
            als.append(_project(truth, sorted(rows)))
        lens = truth.sum(axis=1)
        got = capi.align_path_merge(als, lens)
        assert np.array_equal(got, _oracle_bits(als, n)), trial
        for rows, b in als:
            assert np.array_equal(_project(got, rows)[1], b)


def test_merge_edge_cases():
    # one alignment: returned as it is; rows nobody mentions stay empty
    b = np.array([[1, 0, 1], [1, 1, 0]], np.uint8)
    got = capi.align_path_merge([([0, 2], b)], [2, 5, 2])
    assert np.array_equal(got[[0, 2]], b) and not got[1].any()
    # no alignments at all
    assert capi.align_path_merge([], [3]).shape == (1, 0)
    # residue counts that disagree (alignpath.cpp:113)
    with pytest.raises(capi.HistError) as e:
        capi.align_path_merge([([0, 1], b), ([1, 2], np.array([[1], [1]], np.uint8))], [2, 2, 1])
    assert e.value.code == -1 and "residues" in str(e.value)
    # an empty column (alignpath.cpp:136)
    with pytest.raises(capi.HistError):
        capi.align_path_merge([([0, 1], np.array([[1, 0], [1, 0]], np.uint8))], [1, 1])
    # inconsistent: 0~1 and 1~2 aligned, but 0 and 2 aligned crosswise
    x = np.array([[1, 1], [1, 1]], np.uint8)
    cross = np.array([[0, 1, 1], [1, 1, 0]], np.uint8)
    with pytest.raises(capi.HistError) as e:
        capi.align_path_merge([([0, 1], x), ([1, 2], x), ([0, 2], cross)], [2, 2, 2])
    assert e.value.code == -4
    with pytest.raises(AssertionError):
        oracle_merge(_as_paths([([0, 1], x), ([1, 2], x), ([0, 2], cross)]))


def _tokens(seqs, alphabet):
    lut = np.full(256, -1, np.int8)
    for k, c in enumerate(alphabet):
        lut[ord(c)] = lut[ord(c.upper())] = k
    return [lut[np.frombuffer(s.encode(), np.uint8)] for s in seqs]


def _params(d):
    mode = _text(d["mode"])
    kw = dict(all_pairs=_text(d["graph"]) == "all", seed=int(d["seed"][0]))
    if mode == "full":
        kw["sparse"] = False
    else:
        band, klen, thr, size = (int(v) for v in mode.split(":")[1:])
        kw.update(sparse=True, band_size=band, kmer_len=klen, kmer_threshold=thr, max_size=size)
    return kw


@pytest.mark.gpu
@pytest.mark.parametrize("path", FIX, ids=IDS)
def test_device_guide_alignment_equals_the_reference(ctx, path):
    d = H.load_golden(path)
    alphabet = _text(d["alphabet"])
    A = len(alphabet)
    seqs = _text(d["seqs"]).split("\n")
    bits, info = ctx.guide_align(_tokens(seqs, alphabet), d["submat"].reshape(A, A), d["scores"], **_params(d))
    # the edges, in the fixture's (source, destination) order, with bit-equal scores
    order = np.lexsort((info["edges"][:, 1], info["edges"][:, 0]))
    assert np.array_equal(info["edges"][order], d["edges"])
    assert np.array_equal(info["edge_lp"][order], d["edge_lp"])
    want = _bits(_text(d["gapped"]).split("\n"))
    assert bits.shape == want.shape and np.array_equal(bits, want)


@pytest.mark.gpu
def test_single_sequence_and_pair(ctx):
    d = H.load_golden(os.path.join(H.GOLDEN_DIR, "guide_pf16593_rnd.npz"))
    alphabet = _text(d["alphabet"])
    A = len(alphabet)
    toks = _tokens(_text(d["seqs"]).split("\n")[:2], alphabet)
    sub = d["submat"].reshape(A, A)
    bits, info = ctx.guide_align(toks[:1], sub, d["scores"])
    assert bits.shape == (1, len(toks[0])) and bits.all() and len(info["edges"]) == 0
    bits, info = ctx.guide_align(toks, sub, d["scores"])
    _, _, paths = ctx.quickalign([(toks[0], toks[1], np.arange(1 - len(toks[1]), len(toks[0]), dtype=np.int32))], sub, d["scores"])
    assert np.array_equal(bits[0], paths[0] & 1) and np.array_equal(bits[1], (paths[0] >> 1) & 1)
    assert info["edges"].tolist() == [[0, 1]] and info["mst"].tolist() == [0]


@pytest.mark.parametrize("name", ["guide_pf16593_rnd", "guide_pf16593_all", "guide_lg24x200_rnd_seed7"])
def test_oracle_pipeline_equals_the_reference(name):
    """Pins the restated merge (and the oracle's Viterbi) on the reference's output: the fixture's edges, aligned by the
    oracle inside full envelopes, joined greedily from sequence 0 (span.cpp:119-144), merged by the restated loops, give the
    reference's guide alignment; the library's merge of the same alignments too."""
    d = H.load_golden(os.path.join(H.GOLDEN_DIR, name + ".npz"))
    alphabet = _text(d["alphabet"])
    A = len(alphabet)
    toks = _tokens(_text(d["seqs"]).split("\n"), alphabet)
    sub = d["submat"].reshape(A, A)
    n = len(toks)
    in_tree, als = {0}, []
    while len(in_tree) < n:
        cand = [k for k, (a, b) in enumerate(d["edges"]) if (a in in_tree) != (b in in_tree)]
        k = max(cand, key=lambda q: d["edge_lp"][q])
        assert sum(d["edge_lp"][q] == d["edge_lp"][k] for q in cand) == 1   # no ties: the queue order does not matter here
        a, b = (int(v) for v in d["edges"][k])
        lp, _, path = O.quickalign(toks[a], toks[b], np.arange(1 - len(toks[b]), len(toks[a]), dtype=np.int32), sub, d["scores"])
        assert lp == d["edge_lp"][k]
        als.append(([a, b], np.stack([path & 1, (path >> 1) & 1]).astype(np.uint8)))
        in_tree |= {a, b}
    want = _bits(_text(d["gapped"]).split("\n"))
    assert np.array_equal(_oracle_bits(als, n), want)
    assert np.array_equal(capi.align_path_merge(als, [len(t) for t in toks]), want)
