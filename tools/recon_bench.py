"""Wall time of the progressive reconstruction (BASELINE metric 2, 'recon wall-time') on a synthetic LG family:
historian_b200.recon.Reconstructor on cuda:0 vs the unmodified reference binary (oracle/_ref/bin/historian,
one core) with the same guide alignment, tree, band and deterministic profiles (-profmaxstates 1).
    python tools/recon_bench.py N_LEAVES ROOT_LEN [--ref]"""
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
rec = R.Reconstructor(ctx, lg, band=20, max_profile_states=1)
t0 = time.perf_counter()
path, lp, _ = rec.reconstruct(tree, seqs, grows)
t1 = time.perf_counter()
cells = sum(e["cells"] for e in rec.log)
print("ours: %d leaves x %d: %.2f s wall, %d merges, %.1f M Forward cells, lp %.3f" % (n_leaves, root_len, t1 - t0, len(rec.log), cells / 1e6, lp))
if "--ref" in sys.argv:
    ref = os.path.join(ROOT, "oracle", "_ref", "bin", "historian")
    with tempfile.TemporaryDirectory() as tmp:
        tf, gf = os.path.join(tmp, "tree.nh"), os.path.join(tmp, "guide.fa")
        open(tf, "w").write(tree.newick())
        with open(gf, "w") as f:
            for v, row in zip(leaves, guide):
                f.write(">%s\n%s\n" % (tree.names[v], bytes(row).decode()))
        t0 = time.perf_counter()
        out = subprocess.run([ref, "recon", "-guide", gf, "-tree", tf, "-preset", "lg", "-profmaxstates", "1", "-output", "fasta", "-norefine", "-band", "20"],
                             stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True, cwd=tmp).stdout.decode()
        t1 = time.perf_counter()
    rows = []
    for line in out.splitlines():          # FASTA rows may be wrapped
        if line.startswith(">"):
            rows.append("")
        elif rows:
            rows[-1] += line.strip()
    ours = R.gapped_rows(path, tree, seqs, wildcard="x")
    same = all(ours[k] == r.lower() or ours[k] == r for k, r in enumerate(rows)) and len(rows) == tree.n_nodes
    print("reference binary (1 core): %.2f s wall; identical alignment: %s" % (t1 - t0, same))
    if not same:
        diff_rows = [k for k, r in enumerate(rows) if not (ours[k] == r.lower() or ours[k] == r)]
        print("rows that differ: %d of %d; alignment columns ours %d / reference %d" % (len(diff_rows), len(rows), len(ours[0]), len(rows[0])))
        # residue-level agreement: for every leaf pair (a, b) sample, are the same residues aligned?
        def colmap(rowstr):
            pos, m = 0, {}
            for c, ch in enumerate(rowstr):
                if ch != "-":
                    m[pos] = c; pos += 1
            return m
        lv = [k for k in range(len(rows)) if tree.is_leaf(k)]
        agree = tot = 0
        for a, b in zip(lv[:-1], lv[1:]):
            for rs, acc in ((ours, 0), (rows, 1)):
                pass
            ma, mb = colmap(ours[a]), colmap(ours[b]); ra, rb = colmap(rows[a]), colmap(rows[b])
            inv_o = {c: p for p, c in mb.items()}; inv_r = {c: p for p, c in rb.items()}
            for p in ma:
                tot += 1
                agree += inv_o.get(ma[p], -1) == inv_r.get(ra[p], -1)
        print("adjacent-leaf residue pairings identical: %d of %d (%.4f%%)" % (agree, tot, 100.0 * agree / max(1, tot)))
        for k in diff_rows[:3]:
            d = [c for c in range(min(len(ours[k]), len(rows[k]))) if ours[k][c] != rows[k][c].lower()]
            print("row %d (%s): first differing column %s of %d" % (k, "leaf" if tree.is_leaf(k) else "internal", d[:1], len(d)))
