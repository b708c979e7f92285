#!/usr/bin/env python3
"""Round-2 fixtures from the REAL reference (oracle/_ref/bin/ref_dump): BASELINE's own configs and a full-size
internal-node merge.  Large matrices are dumped sparsely (REF_DUMP_SPARSE=<stride>, see oracle/ref_dump.cpp):
every stride-th in-envelope cell plus the per-row envelope counts, so the fixtures stay small.

  gp120_nodes.npz      config 1: data/gp120.fa with its guide and tree, root + two internal nodes (band 20, sampled profiles)
  ecmrest8.npz         config 5: an ECMrest codon family (61-state), 8 leaves x 60 codons, every node (dense)
  lg16x300_n14.npz     a depth-3 node of a 16-leaf x 300-aa LG family at band 20, -profsamples 40 -profmaxstates 4000:
                       DAG children (in-degree > 1), anti-diagonals wider than 32 and 96 slots
  pf16593/             config 2: PF16593.fa, its tree and the reference's own guide alignment of it (data/PF16593.guide.fa),
                       for the whole-tree sampled run against oracle/_ref/bin/historian at test time
"""
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import make_golden as G
from historian_b200 import descs as D

REF, BIN, OUT = G.REF, G.BIN, G.OUT


def dump(name, mode, recon_args, which, sparse=0, drop=()):
    with tempfile.TemporaryDirectory() as tmp:
        binf = os.path.join(tmp, "d.bin")
        env = dict(os.environ)
        if sparse:
            env["REF_DUMP_SPARSE"] = str(sparse)
        subprocess.run([os.path.join(BIN, "ref_dump"), binf, which, mode] + recon_args, check=True, cwd=REF, env=env,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        d = D.read_dump(binf)
    d = {k: (np.frombuffer(v.encode(), np.uint8) if isinstance(v, str) else v) for k, v in d.items()
         if not any(k.endswith(s) for s in drop)}
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **d)
    print(name, "nodes", list(d["dumped_nodes"]), "%.1f kB" % (os.path.getsize(path) / 1e3))


# arrays the round-2 tests do not read (list orders are pinned by the round-1 fixtures)
DROP = ("in_flat", "in_off", "null_flat", "null_off", "abs_flat", "abs_off", "cum_lp6", "post_lp", "post_cells", "post_threshold")


def main():
    # ---- config 1: gp120 (10 leaves -> nodes 0..18, root 18)
    dump("gp120_nodes", "sample", ["-guide", "data/gp120.guide.fa", "-tree", "data/gp120.tree.nh", "-band", "20", "-profsamples", "10",
                                   "-profmaxstates", "1500", "-seed", "5"], which="auto3", sparse=11, drop=DROP)
    with tempfile.TemporaryDirectory() as tmp:
        # ---- config 5: ECMrest codon family (the reference simulates and aligns in its one-character codon tokens)
        tf, gf = G.simulate(tmp, G.balanced(8, 0.2), ["-preset", "ECMrest", "-rootlen", "60", "-seed", "4"], 8)
        dump("ecmrest8", "sample", ["-guide", gf, "-tree", tf, "-preset", "ECMrest", "-band", "8", "-profsamples", "4", "-profmaxstates", "250",
                                    "-seed", "3"], which="all")
        # ---- a full-size internal merge
        # (indel rates and branch lengths chosen so that the sampled traces really differ: ~1000-state DAG children,
        # ~300 in-envelope cells per row, a 3300-state parent)
        tf2, gf2 = G.simulate(tmp, G.balanced(16, 0.5), ["-preset", "lg", "-rootlen", "300", "-seed", "2", "-insrate", "0.1", "-delrate", "0.1"], 16)
        dump("lg16x300_n14", "sample", ["-guide", gf2, "-tree", tf2, "-preset", "lg", "-band", "20", "-profsamples", "40", "-profmaxstates", "4000",
                                        "-seed", "9"], which="14", sparse=47, drop=DROP + (".subx", ".suby"))
    # ---- config 2: inputs of the whole-tree run
    pf = os.path.join(OUT, "pf16593")
    os.makedirs(pf, exist_ok=True)
    for f in ("PF16593.fa", "PF16593.nhx", "PF16593.guide.fa"):
        shutil.copy(os.path.join(REF, "data", f), os.path.join(pf, f))
        os.chmod(os.path.join(pf, f), 0o644)
    # the reference's own Stockholm output of that run (its #=GF LP line is what tests/test_loglike.py reproduces)
    import subprocess
    hist = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "bin", "historian")
    for name, extra in (("PF16593.recon.stk", []), ("PF16593.recon.ancseq.stk", ["-ancseq"])):
        txt = subprocess.run([hist, "recon", "-seqs", os.path.join(pf, "PF16593.fa"), "-tree", os.path.join(pf, "PF16593.nhx"), "-output",
                              "stockholm"] + extra, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, check=True).stdout
        txt = "\n".join("#=GF ID             PF16593.fa" if l.startswith("#=GF ID") else l for l in txt.split("\n"))
        open(os.path.join(pf, name), "w").write(txt)
    print("pf16593:", os.listdir(pf))


if __name__ == "__main__":
    main()
