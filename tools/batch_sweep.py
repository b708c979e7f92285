"""Resident Forward+Backward fill throughput against the number of leaf x leaf merges per launch:
families of 2 * pairs leaves of the bench's shape, band 20.  Prints one line per batch size."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from historian_b200 import capi

ctx = capi.Context(0)
stream = torch.cuda.ExternalStream(ctx.stream, device=0)
for n_pairs in [int(a) for a in sys.argv[1:]] or [128, 256, 512, 1024, 2048, 4096]:
    lg, guide = bench.make_family(1, n_leaves=2 * n_pairs)
    hmm, inputs = bench.cherry_inputs(lg, guide, n_pairs)
    leaves = ctx.upload_sequences(bench.cherry_sequences(guide, n_pairs), lg.alphabet, lg.components)
    pairs = ctx.prepare_batch(leaves[0::2], leaves[1::2], hmm, [e for _, _, e in inputs], backward=True)
    cells = 2 * sum(p.n_cells for p in pairs)
    for _ in range(3):
        ctx.fill_batch(pairs, backward=True)
    handles = ctx.fill_batch_async(pairs, backward=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    steps = 10
    for _ in range(steps):
        ctx.fill_batch_async(pairs, backward=True, arr=handles)
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    print("pairs %5d  strips %5d  cells/launch %.1f M  %.3f ms/launch  %.1f G cells/s" % (n_pairs, 2 * n_pairs, cells / 1e6, ms, cells / ms / 1e6), flush=True)
    ctx.close_all(pairs)
    ctx.close_all(leaves)
