"""Small end-to-end loop over the batched DPMatrix constructors (debugging aid; HIST_B200_VERIFY_PREP=1 compares the
device-built band keys and layout with a host computation)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from historian_b200 import capi
n_pairs = int(sys.argv[1]) if len(sys.argv) > 1 else 8
lg, guide = bench.make_family(1, n_leaves=2 * n_pairs)
hmm, inputs = bench.cherry_inputs(lg, guide, n_pairs)
ctx = capi.Context(0)
seqs = bench.cherry_sequences(guide, n_pairs)
plan = None
first = None
for it in range(3):
    lv = ctx.upload_sequences(seqs, lg.alphabet, lg.components)
    if plan is None:
        plan = capi.PrepareBatch(ctx, lv[0::2], lv[1::2], hmm, [e for _, _, e in inputs], backward=True)
    pr = plan.run(lv[0::2], lv[1::2])
    r = ctx.fill_batch(pr, backward=True)
    print(it, r[0][:4], r[1][:4], "finite:", int(np.isfinite(r[0]).sum()), "of", n_pairs, "h2d", ctx.h2d_bytes)
    if first is None:
        first = r
    assert np.array_equal(first[0], r[0]) and np.all(np.isfinite(r[0])) and np.allclose(r[0], r[1], rtol=1e-5)
    ctx.close_all(pr)
    ctx.close_all(lv)
print("ok")
