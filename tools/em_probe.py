"""Kernel time of the sum-product column kernel on dense vs sparse column subsets of the EM bench alignment."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from historian_b200 import capi
model, tree, cols = bench.em_family()
ctx = capi.Context(0)
sp = ctx.sumprod(model, tree)
occ = (cols != ord("-")).sum(1)
print("columns", cols.shape, "dense(>N/2):", int((occ > cols.shape[1] / 2).sum()), "median occ", int(np.median(occ)), "mean", float(occ.mean()))
def run(sub, label):
    d = torch.from_numpy(np.ascontiguousarray(sub)).cuda()
    c = torch.zeros(sp.counts_size, dtype=torch.float64, device="cuda")
    ll = torch.zeros(len(sub), dtype=torch.float64, device="cuda")
    for _ in range(2): sp.columns_device(d, len(sub), c, ll)
    ctx.set_timing(True); ctx.fill_time_ms()
    for _ in range(3): sp.columns_device(d, len(sub), c, ll)
    ms, n = ctx.fill_time_ms(); ctx.set_timing(False)
    print("%-28s %6d cols  %.3f ms  (%.1f us/col)" % (label, len(sub), ms / n, 1e3 * ms / n / len(sub)))
run(cols, "all")
run(cols[occ > cols.shape[1] / 2], "dense")
run(cols[occ <= cols.shape[1] / 2], "sparse")
run(cols[occ <= 8], "very sparse (<=8 rows)")
run(cols[:888], "first 888")
