"""Wall time of the native level-batched reconstruction (BASELINE metric 2) on a synthetic LG family, band 20.
    python tools/recon_native_bench.py N_LEAVES ROOT_LEN [--samples] [--py] [--ref]"""
import os, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from historian_b200 import capi, model as M, recon as R, simulator as S

n_leaves, root_len = int(sys.argv[1]), int(sys.argv[2])
tree = S.Tree.balanced(n_leaves, 0.05)
lg = M.RateModel.preset("lg")
gapped, _ = S.simulate_family(lg, tree, root_len, 1)
leaves, guide = S.leaf_guide(tree, gapped)
seqs = {v: bytes(row[row != ord("-")]).decode() for v, row in zip(leaves, guide)}
grows = {v: row != ord("-") for v, row in zip(leaves, guide)}
ctx = capi.Context(0)
sampled = "--samples" in sys.argv
kw = dict(band=20, profile_samples=10, max_profile_states=5000 if sampled else 1, seed=1)
for rep in range(2):
    nat = R.NativeReconstructor(ctx, lg, **kw)
    t0 = time.perf_counter()
    bits, lp, rec = nat.reconstruct(tree, seqs, grows)
    t1 = time.perf_counter()
    print("native (%s): %d leaves x %d: %.3f s wall, %d columns, lp %.3f" % ("sampled" if sampled else "best-trace", n_leaves, root_len, t1 - t0, bits.shape[1], lp))
    print("  stages ms: " + ", ".join("%s %.1f" % kv for kv in nat.stage_ms.items()))
    rec.close()
if "--py" in sys.argv:
    py = R.Reconstructor(ctx, lg, **kw)
    t0 = time.perf_counter()
    path, lp2, _ = py.reconstruct(tree, seqs, grows)
    t1 = time.perf_counter()
    ncol = max(len(v) for v in path.values())
    want = np.zeros((tree.n_nodes, ncol), bool)
    for row, rp in path.items():
        want[row, :len(rp)] = rp
    print("python driver: %.3f s wall; identical alignment: %s; lp equal: %s" % (t1 - t0, want.shape == bits.shape and bool((want == bits).all()), lp2 == lp))
if "--ref" in sys.argv:
    ref = os.path.join(ROOT, "oracle", "_ref", "bin", "historian")
    with tempfile.TemporaryDirectory() as tmp:
        tf, gf = os.path.join(tmp, "tree.nh"), os.path.join(tmp, "guide.fa")
        open(tf, "w").write(tree.newick())
        with open(gf, "w") as f:
            for v, row in zip(leaves, guide):
                f.write(">%s\n%s\n" % (tree.names[v], bytes(row).decode()))
        t0 = time.perf_counter()
        args = [ref, "recon", "-guide", gf, "-tree", tf, "-preset", "lg", "-output", "fasta", "-norefine", "-band", "20", "-seed", "1"]
        args += ["-profsamples", "10", "-profmaxstates", "5000"] if sampled else ["-profmaxstates", "1"]
        out = subprocess.run(args, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True, cwd=tmp).stdout.decode()
        t1 = time.perf_counter()
    rows = []
    for line in out.splitlines():
        if line.startswith(">"):
            rows.append("")
        elif rows:
            rows[-1] += line.strip()
    ours = R.gapped_rows_dense(bits, tree, seqs, wildcard="x")
    same = len(rows) == tree.n_nodes and all(ours[k] == r.lower() or ours[k] == r for k, r in enumerate(rows))
    print("reference binary (1 core): %.2f s wall; identical alignment: %s" % (t1 - t0, same))
