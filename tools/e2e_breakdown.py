"""Where the end-to-end step of bench.py spends its time (host clock around each C-ABI stage)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from historian_b200 import capi

lg, guide = bench.make_family(1)
hmm, inputs = bench.cherry_inputs(lg, guide, 512)
ctx = capi.Context(0)
seqs = bench.cherry_sequences(guide, 512)
plan = None
for it in range(3):
    t0 = time.perf_counter()
    lv = ctx.upload_sequences(seqs, lg.alphabet, lg.components)
    pp = list(zip(lv[0::2], lv[1::2]))
    t1 = time.perf_counter()
    if plan is None:
        plan = capi.PrepareBatch(ctx, lv[0::2], lv[1::2], hmm, [e for _, _, e in inputs], backward=True)
    pr = plan.run(lv[0::2], lv[1::2])
    ctx.synchronize()
    t2 = time.perf_counter()
    r = ctx.fill_batch(pr, backward=True)
    t3 = time.perf_counter()
    ctx.close_all(pr)
    ctx.close_all(lv)
    t4 = time.perf_counter()
    print("upload %.1f ms  prepare %.1f ms  fill %.1f ms  close %.1f ms" % (1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2), 1e3 * (t4 - t3)))
