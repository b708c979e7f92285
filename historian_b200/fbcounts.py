"""BackwardMatrix::getCounts (reference forward.cpp:1183-1214): Forward-Backward expected counts of the root
merge, the E-step of `historian fit` on unaligned sequences (recon.cpp:1012-1013).

The numeric work runs on the device: hist_fb_counts reduces the filled Forward and Backward matrices to six
indel counts and to posterior weights of alignment columns; the weighted columns then go through the same
hist_sumprod_columns batch the aligned E-step uses (em.py).  This module only assembles the column bytes,
which is what ForwardMatrix::getAlignmentColumn (forward.cpp:940-973) does one cell at a time.

Scope: profile transitions are taken to carry no counts of their own (see hist_fb_counts), so the result is
the reference's getCounts() for a merge of two leaf profiles, or of profiles built without
CountSubstEvents / CountIndelEvents."""
import numpy as np

from . import capi

GAP, WILD = ord("-"), ord("*")


def leaf_columns(seq, row, n_rows):
    """Profile::alignColumn (profile.cpp:100-110) of a leaf profile: state i (1..L) holds residue i-1 in `row`."""
    seq = np.frombuffer(seq.encode() if isinstance(seq, str) else bytes(seq), np.uint8)
    cols = np.full((len(seq) + 2, n_rows), GAP, np.uint8)
    cols[1:len(seq) + 1, row] = seq
    return cols


def prob_model_waits(l_probs, r_probs):
    """The constants transitionEigenCounts adds (forward.cpp:579-652), from the two model.ProbModel objects."""
    return np.array([l_probs.t, r_probs.t, l_probs.insWait, l_probs.delWait, r_probs.insWait, r_probs.delWait])


def column_batch(x_cols, y_cols, parent_row, w_x, w_y, cells, cell_w):
    """Weighted column batch of getCounts: ([n, n_rows] uint8 columns, [n] weights)."""
    x_cols, y_cols = np.asarray(x_cols, np.uint8), np.asarray(y_cols, np.uint8)
    cols, weights = [], []
    for side, w in ((x_cols, w_x), (y_cols, w_y)):
        nonempty = (side != GAP).any(axis=1)
        keep = np.flatnonzero((w[0] > 0) & nonempty)          # insert columns: the child's own column
        cols.append(side[keep])
        weights.append(w[0][keep])
        keep = np.flatnonzero((w[1] > 0) & nonempty)          # deletion columns: plus a wildcard parent
        c = side[keep].copy()
        c[:, parent_row] = WILD
        cols.append(c)
        weights.append(w[1][keep])
    if len(cells):                                            # match columns: both children and the parent
        cx, cy = x_cols[cells[:, 0]], y_cols[cells[:, 1]]
        c = np.where(cx != GAP, cx, cy)
        c[:, parent_row] = WILD
        cols.append(c)
        weights.append(np.asarray(cell_w))
    return np.concatenate(cols), np.concatenate(weights)


def fb_counts(pair, sumprod, waits, x_cols, y_cols, parent_row, lp_end):
    """getCounts of a pair whose Forward and Backward matrices are filled (`lp_end` = the Forward likelihood).
    Returns the flat EigenCounts buffer of `sumprod` (indel[7] | rootCount | eigenCount; SumProd.transform_counts turns it into EventCounts)."""
    indel, w_x, w_y, cells, cell_w = pair.fb_counts(waits)
    cols, weights = column_batch(x_cols, y_cols, parent_row, w_x, w_y, cells, cell_w)
    counts = np.zeros(sumprod.counts_size)
    if len(cols):
        sumprod.columns(cols, weights=weights, flags=capi.SP_COUNTS, counts=counts, col_loglike=False)
    counts[:6] = indel
    counts[6] = lp_end
    return counts
