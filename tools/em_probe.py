"""One E-step of the EM leg (bench.py's em_family) for kernel-level timing under ncu:
    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv python tools/em_probe.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench
from historian_b200 import capi
model, tree, cols = bench.em_family()
ctx = capi.Context(0)
sp = ctx.sumprod(model, tree)
dev = torch.device("cuda", 0)
d_cols = torch.from_numpy(np.ascontiguousarray(cols)).to(dev)
d_counts = torch.zeros(sp.counts_size, dtype=torch.float64, device=dev)
d_ll = torch.zeros(len(cols), dtype=torch.float64, device=dev)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    d_counts.zero_()
    torch.cuda.synchronize()
    sp.columns_device(d_cols, len(cols), d_counts, d_ll)
    ctx.synchronize()
print("lp", float(d_counts[6]))
