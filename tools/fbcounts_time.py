"""Wall time of BackwardMatrix::getCounts on the device (hist_fb_counts + the weighted sum-product batch) for one
leaf x leaf merge of the bench family (1000 residues, LG, band 20)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import bench
from historian_b200 import capi, descs as D, fbcounts, model as M

lg, guide = bench.make_family(1)
hmm, inputs = bench.cherry_inputs(lg, guide, 1)
seqs = bench.cherry_sequences(guide, 1)
ctx = capi.Context(0)
x, y = ctx.upload_sequences(seqs, lg.alphabet, lg.components)
pair = ctx.prepare(x, y, hmm, inputs[0][2])
lp_end = pair.forward()
pair.backward()
t = bench.BRANCH
sp = ctx.sumprod(D.SumProdModel.from_rate_model(lg), D.TreeTopology([2, 2, -1], [t, t, 0.0]))
waits = fbcounts.prob_model_waits(M.ProbModel(lg, t), M.ProbModel(lg, t))
xc, yc = fbcounts.leaf_columns(seqs[0], 0, 3), fbcounts.leaf_columns(seqs[1], 1, 3)
for it in range(3):
    t0 = time.perf_counter()
    raw = pair.fb_counts(waits)
    t1 = time.perf_counter()
    cols, w = fbcounts.column_batch(xc, yc, 2, raw[1], raw[2], raw[3], raw[4])
    t2 = time.perf_counter()
    counts = np.zeros(sp.counts_size)
    sp.columns(cols, weights=w, flags=capi.SP_COUNTS, counts=counts, col_loglike=False)
    t3 = time.perf_counter()
    print("cells %d  match columns %d  hist_fb_counts %.2f ms  column assembly %.2f ms  sum-product %.2f ms  indel %s" % (
        pair.n_cells, len(raw[3]), 1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2), np.round(raw[0], 3)))
