"""Host side of the guide-alignment stage's pairwise step: what the QuickAlignMatrix constructor
derives from (model, time) before its fill (quickalign.cpp:22-50) and the k-mer seeded DiagonalEnvelope with a fixed
threshold (diagenv.cpp:101-190, the kmerThreshold >= 0 branch).  The Viterbi fill and traceback run on the device
(hist_quickalign_batch, csrc/hist_quickalign.cu)."""
import math

import numpy as np

from .model import ProbModel


def quickalign_scores(model, time):
    """(submat [A][A], (noGap, gapOpen, gapExtend, m2m, m2i, m2d, i2i, i2m, i2d, d2d, d2m)), quickalign.cpp:22-50."""
    pm = ProbModel(model, time)
    with np.errstate(divide="ignore"):
        sub = np.log(np.asarray(pm.subMat[0], float)) - np.log(np.asarray(pm.insVec[0], float))[None, :]
    gap = pm.ins + (1 - pm.ins) * pm.del_
    no_gap = 1 - gap
    gap_ext = 1 / ((pm.ins / gap) / pm.insExt + (1 - pm.ins / gap) / pm.delExt)
    no_gap_ext = 1 - gap_ext
    lg = math.log
    return sub, (lg(no_gap), lg(gap) + lg(no_gap_ext), lg(gap_ext), lg(no_gap * no_gap), lg(gap), lg(no_gap * gap),
                 lg(gap_ext), lg(no_gap_ext * no_gap), lg(no_gap_ext * gap), lg(gap_ext), lg(no_gap_ext))


def sparse_envelope(x_tok, y_tok, alph_size, kmer_len=6, band=64, threshold=14):
    """DiagonalEnvelope::initSparse with kmerThreshold >= 0 (diagenv.cpp:101-190): diagonals i - j (0-based k-mer starts)
    with at least `threshold` shared k-mers seed bands of +-band/2; diagonal 0 is always present; sequences shorter than
    2 * (kmer_len + threshold) get the full envelope.  Returns the ascending diagonal list."""
    x, y = np.asarray(x_tok, np.int64), np.asarray(y_tok, np.int64)
    lx, ly = len(x), len(y)
    lo, hi = 1 - ly, lx - 1
    if lx < 2 * (kmer_len + threshold) or ly < 2 * (kmer_len + threshold):
        return np.arange(lo, hi + 1, dtype=np.int32)

    def kmers(t):
        n = len(t) - kmer_len + 1
        w = np.lib.stride_tricks.sliding_window_view(t, kmer_len)
        valid = np.all(w >= 0, axis=1)
        code = (np.where(w < 0, 0, w) * (alph_size ** np.arange(kmer_len - 1, -1, -1, dtype=np.int64))).sum(axis=1)   # makeKmer, fastseq.cpp
        return np.arange(n)[valid], code[valid]

    xi, xc = kmers(x)
    yj, yc = kmers(y)
    order = np.argsort(yc, kind="stable")
    ys, yjs = yc[order], yj[order]
    left, right = np.searchsorted(ys, xc, "left"), np.searchsorted(ys, xc, "right")
    counts = {}
    for i, a, b in zip(xi[right > left], left[right > left], right[right > left]):
        for j in yjs[a:b]:
            counts[i - j] = counts.get(i - j, 0) + 1
    diags = {0}
    half = band // 2
    for d, c in counts.items():
        if c >= threshold:
            diags.update(range(max(lo, d - half), min(hi, d + half) + 1))
    return np.array(sorted(diags), np.int32)


def envelope_cells(diagonals, x_len, y_len):
    """Cells the fill visits: for every diagonal the (i, j) with i - j = d, 1 <= i <= xLen, 1 <= j <= yLen."""
    d = np.asarray(diagonals, np.int64)
    return int(np.maximum(0, np.minimum(y_len, x_len - d) - np.maximum(1, 1 - d) + 1).sum())


def expected_substitution_rate(model):
    """RateModel::expectedSubstitutionRate (model.cpp:336-347): sum_c w_c sum_{i != j} eqm_c[i] R_c[i][j]."""
    rate = 0.0
    for c in range(model.components):
        r = np.asarray(model.subRate[c], float)
        eqm = np.asarray(model.getEqmProbVector(r), float)
        rate += float(model.cptWeight[c]) * float((eqm[:, None] * r).sum() - (eqm * np.diag(r)).sum())
    return rate


def distance_matrix(ctx, model, rows, max_iterations=100):
    """RateModel::distanceMatrix (model.cpp:533-553) of a gapped alignment (uint8 [n][cols]) on the device."""
    from . import descs as D
    return ctx.distance_matrix(D.SumProdModel.from_rate_model(model), expected_substitution_rate(model), rows, max_iterations)


def newick(names, child, branch_length):
    """Tree::toString (tree.cpp:56-102) of the tree hist_build_tree returns: internal nodes unnamed, branch lengths with the
    stream's default 6 significant digits -- the text the reference parses back (tree.cpp:347-350), so its trees carry these
    rounded lengths."""
    import sys
    n = len(names)
    sys.setrecursionlimit(max(sys.getrecursionlimit(), 4 * n + 100))

    def rec(k):
        if child[k][0] < 0:
            return names[k]
        return "(" + ",".join("%s:%s" % (rec(int(c)), "%g" % branch_length[int(c)]) for c in child[k]) + ")"
    return rec(2 * n - 2) + ";"


def build_tree(ctx, names, dist, method="upgma"):
    """Tree::buildByUPGMA / buildByNeighborJoining on the device -> Newick text equal to the reference's."""
    child, blen = ctx.build_tree(dist, 0 if method == "upgma" else 1)
    return newick(list(names), child, blen)


def tokenize(seqs, alphabet):
    """FastSeq::unvalidatedTokens (fastseq.cpp): token of every character, case-insensitive, -1 outside the alphabet."""
    lut = np.full(256, -1, np.int8)
    for k, c in enumerate(alphabet):
        lut[ord(c.lower())] = lut[ord(c.upper())] = k
    return [lut[np.frombuffer(s.encode(), np.uint8)] for s in seqs]


def guide_alignment(ctx, model, seqs, time=1.0, **params):
    """AlignGraph::mstAlign as `historian recon -seqs` runs it (recon.cpp:806-821, time 1): hist_guide_align on the device.
    params: all_pairs, seed, sparse, kmer_len, kmer_threshold, band_size, max_size (Context.guide_align).
    Returns (bits uint8 [n][cols], info)."""
    sub, sc = quickalign_scores(model, time)
    return ctx.guide_align(tokenize(seqs, model.alphabet), sub, sc, **params)


def gapped(bits, seqs, gap="-"):
    """Alignment::gapped (alignpath.cpp:245-276)."""
    rows = []
    for b, s in zip(bits, seqs):
        r = np.full(len(b), ord(gap), np.uint8)
        r[np.flatnonzero(b)] = np.frombuffer(s.encode(), np.uint8)
        rows.append(r.tobytes().decode())
    return rows


def tree_from_sequences(ctx, model, names, seqs, method="upgma", max_iterations=100, **guide_params):
    """The front end of `historian recon -seqs` (recon.cpp:806-821, 731-743): guide alignment, distance matrix of its gapped
    rows, UPGMA (the default) or neighbour-joining tree -- every stage on the device.  Returns (guide bits [n][cols], Newick)."""
    bits, _ = guide_alignment(ctx, model, seqs, **guide_params)
    rows = np.array([np.frombuffer(r.encode(), np.uint8) for r in gapped(bits, seqs)])
    dist = distance_matrix(ctx, model, rows, max_iterations)
    return bits, build_tree(ctx, names, dist, method)
