"""Widest anti-diagonal (in stored cells) of the leaf x leaf merges of a bench family: which merges miss the chain kernel."""
import sys, os, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from historian_b200 import capi
ctx = capi.Context(0)
for seed in [int(a) for a in sys.argv[1:]] or [1, 3, 8]:
    lg, guide = bench.make_family(seed)
    hmm, inputs = bench.cherry_inputs(lg, guide, 512)
    lv = ctx.upload_sequences(bench.cherry_sequences(guide, 512), lg.alphabet, lg.components)
    pr = ctx.prepare_batch(lv[0::2], lv[1::2], hmm, [e for _, _, e in inputs], backward=True)
    w = [int(p.info.max_diag_width) for p in pr]
    h = collections.Counter(w)
    print("seed", seed, "max width", max(w), "pairs wider than 30:", [(k, x) for k, x in enumerate(w) if x > 30], "histogram tail", sorted(h.items())[-6:])
    ctx.close_all(pr); ctx.close_all(lv)
