"""Host-clock split of the whole-tree reconstruction's calls (create / set_guide / add_leaves / levels / alignment)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from historian_b200 import capi, recon as R
lg, guide, tree, leaves = bench.make_family(1, with_tree=True)
seqs, grows = bench._family_inputs(tree, leaves, guide)
ctx = capi.Context(0)
for rep in range(3):
    nat = R.NativeReconstructor(ctx, lg, band=20, max_profile_states=1)
    t = [time.perf_counter()]
    rec = capi.Recon(ctx, tree.n_nodes, lg.alphabet, lg.components); t.append(time.perf_counter())
    lv = tree.leaves()
    g = np.array([grows[v] for v in lv], np.uint8); t.append(time.perf_counter())
    rec.set_guide(lv, g); t.append(time.perf_counter())
    rec.add_leaves(lv, [seqs[v] for v in lv]); t.append(time.perf_counter())
    groups = nat.schedule(tree)
    bits, lp, rec = nat.reconstruct(tree, seqs, grows, rec=rec, groups=groups)
    t.append(time.perf_counter())
    names = ["create", "guide array (python)", "set_guide", "add_leaves", "levels + alignment"]
    print(" | ".join("%s %.1f ms" % (n, 1e3 * (b - a)) for n, a, b in zip(names, t, t[1:])), "| stages sum %.1f ms" % sum(nat.stage_ms.values()))
    t0 = time.perf_counter(); rec.alignment(); print("alignment() alone %.1f ms" % (1e3 * (time.perf_counter() - t0)))
    rec.close()

# per level: host clock around the Python call against the library's own stage timers
nat = R.NativeReconstructor(ctx, lg, band=20, max_profile_states=1)
rec = capi.Recon(ctx, tree.n_nodes, lg.alphabet, lg.components)
rec.set_guide(lv, g)
rec.add_leaves(lv, [seqs[v] for v in lv])
orig = rec.level
acc = {"py": 0.0}
prev = [0.0]
def timed(*a, **k):
    t0 = time.perf_counter()
    r = orig(*a, **k)
    dt = 1e3 * (time.perf_counter() - t0)
    st = sum(rec.stage_ms().values())
    print("  level of %4d merges: python call %.1f ms, library stages %.1f ms" % (len(a[0]), dt, st - prev[0]))
    prev[0] = st
    acc["py"] += dt
    return r
rec.level = timed
t0 = time.perf_counter()
bits, lp, rec = nat.reconstruct(tree, seqs, grows, rec=rec, groups=groups)
print("reconstruct %.1f ms, inside level calls %.1f ms" % (1e3 * (time.perf_counter() - t0), acc["py"]))
