#!/usr/bin/env python3
"""Generates tests/golden/sumprod_*.npz from the REAL reference (oracle/_ref/bin/ref_dump_sumprod:
the unmodified AlignColSumProduct / EigenModel compiled against oracle/gslshim).  Run in the build
container where /root/reference exists; the fixtures travel to the GPU box.

Each fixture holds the model's eigen-system, the tree, the per-branch tables, the alignment columns
and the reference's per-column messages, log-likelihoods, posteriors and accumulated counts."""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from historian_b200 import descs as D

REF = os.environ.get("HISTORIAN_REF", "/root/reference")
BIN = os.path.join(ROOT, "oracle", "_ref", "bin")
OUT = os.path.join(ROOT, "tests", "golden")


def named_balanced(n, bl):
    """Newick with every node named (the simulator prints ancestors under these names)."""
    counter = [0]

    def rec(lo, hi):
        if hi - lo == 1:
            return "s%d:%g" % (lo, bl * (1 + 0.3 * (lo % 3)))
        mid = (lo + hi) // 2
        a, b = rec(lo, mid), rec(mid, hi)
        counter[0] += 1
        return "(%s,%s)n%d:%g" % (a, b, counter[0], bl)
    return rec(0, n).rsplit(":", 1)[0] + ";\n"


def read_fasta(text):
    seqs, name = {}, None
    for line in text.splitlines():
        if line.startswith(">"):
            name = line[1:].strip(); seqs[name] = ""
        elif name:
            seqs[name] += line.strip()
    return seqs


def simulate(tmp, tree, args):
    tf = os.path.join(tmp, "tree.nh")
    open(tf, "w").write(tree)
    out = subprocess.run([os.path.join(BIN, "historian"), "generate"] + args + [tf, "-output", "fasta"],
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True).stdout.decode()
    return tf, read_fasta(out)


def dump(name, model, seqs, tf, tmp):
    fa = os.path.join(tmp, name + ".fa")
    with open(fa, "w") as f:
        for k, v in seqs.items():
            f.write(">%s\n%s\n" % (k, v))
    binf = os.path.join(tmp, name + ".bin")
    subprocess.run([os.path.join(BIN, "ref_dump_sumprod"), binf, model, fa, tf], check=True, cwd=REF,
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    d = D.read_dump(binf)
    d = {k: (np.frombuffer(v.encode(), np.uint8) if isinstance(v, str) else v) for k, v in d.items()}
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **d)
    print(name, "C,A,N", list(d["dims"]), "columns", d["columns"].shape[0], "%.1f kB" % (os.path.getsize(path) / 1e3))


def main():
    with tempfile.TemporaryDirectory() as tmp:
        # 1. the reference's own sum-product test inputs (Makefile:293-299)
        dump("sumprod_testaligncount", "data/testnj.jukescantor.json",
             read_fasta(open(os.path.join(REF, "data/testaligncount.fa")).read()), os.path.join(REF, "data/testaligncount.nh"), tmp)
        # 2. DNA, Jukes-Cantor with a 2-component gamma mixture, 12 leaves, full history observed
        #    (the shape of BASELINE config 4: `fit -stockrecon` on a simulated history)
        tf, seqs = simulate(tmp, named_balanced(12, 0.15), ["-preset", "jc", "-gamma", "2", "-shape", "1", "-rootlen", "30", "-seed", "5",
                                                            "-insrate", "0.05", "-delrate", "0.05"])
        dump("sumprod_dna12_gamma2_observed", "preset:jc:2:1", seqs, tf, tmp)
        # 3. same history with the ancestors as wildcards (what `reconstruct -ancseq` / `count -recon` sees)
        wild = {k: (v if k.startswith("s") else "".join("*" if c not in "-." else c for c in v)) for k, v in seqs.items()}
        dump("sumprod_dna12_gamma2_wild", "preset:jc:2:1", wild, tf, tmp)
        # 4. amino acids (LG), 8 leaves, wildcard ancestors, long branches
        tf2, seqs2 = simulate(tmp, named_balanced(8, 0.4), ["-preset", "lg", "-rootlen", "16", "-seed", "9", "-insrate", "0.04", "-delrate", "0.04"])
        wild2 = {k: (v if k.startswith("s") else "".join("*" if c not in "-." else c for c in v)) for k, v in seqs2.items()}
        dump("sumprod_amino8_wild", "preset:lg", wild2, tf2, tmp)


if __name__ == "__main__":
    main()
