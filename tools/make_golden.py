#!/usr/bin/env python3
"""Generates tests/golden/*.npz from the REAL reference (oracle/_ref/bin/ref_dump, i.e.
the unmodified reference sources compiled against oracle/gslshim).  Run in the build
container where /root/reference exists; the fixtures travel to the GPU box.

Each fixture holds, for every internal node of a small progressive reconstruction, the
flattened inputs of the hot path and the reference's outputs (see oracle/ref_dump.cpp)."""
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


def balanced(n, bl, prefix="s"):
    def rec(lo, hi):
        if hi - lo == 1:
            return "%s%d:%g" % (prefix, lo, bl)
        mid = (lo + hi) // 2
        return "(%s,%s):%g" % (rec(lo, mid), rec(mid, hi), bl)
    return rec(0, n).rsplit(":", 1)[0] + ";\n"


def simulate(tmp, tree, args, nleaves):
    """historian generate -> leaf-only gapped guide alignment (FASTA)."""
    tf = os.path.join(tmp, "tree.nh")
    open(tf, "w").write(tree)
    out = subprocess.run([os.path.join(BIN, "historian"), "generate"] + args + [tf, "-output", "fasta"],
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True).stdout.decode()
    seqs, name = {}, None
    for line in out.splitlines():
        if line.startswith(">"):
            name = line[1:].strip(); seqs[name] = ""
        elif name:
            seqs[name] += line.strip()
    leaves = {k: v for k, v in seqs.items() if k.startswith("s") and k[1:].isdigit()}
    assert len(leaves) == nleaves
    L = len(next(iter(leaves.values())))
    cols = [i for i in range(L) if any(v[i] not in "-." for v in leaves.values())]
    gf = os.path.join(tmp, "guide.fa")
    with open(gf, "w") as f:
        for k, v in leaves.items():
            f.write(">%s\n%s\n" % (k, "".join(v[i] for i in cols)))
    return tf, gf


def dump(name, mode, recon_args, which="all"):
    with tempfile.TemporaryDirectory() as tmp:
        binf = os.path.join(tmp, "d.bin")
        args = [a.replace("@TMP", tmp) for a in recon_args]
        subprocess.run([os.path.join(BIN, "ref_dump"), binf, which, mode] + args, check=True, cwd=REF,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        d = D.read_dump(binf)
    d = {k: (np.frombuffer(v.encode(), np.uint8) if isinstance(v, str) else v) for k, v in d.items()}
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
    print(name, "nodes", list(d["dumped_nodes"]), "%.1f kB" % (os.path.getsize(os.path.join(OUT, name + ".npz")) / 1e3))


def main():
    with tempfile.TemporaryDirectory() as tmp:
        # 1. amino (LG), 8 leaves x ~40 residues, band 4, default sampling mode -> DAG profiles
        tf, gf = simulate(tmp, balanced(8, 0.3), ["-preset", "lg", "-rootlen", "40", "-seed", "3", "-insrate", "0.05", "-delrate", "0.05"], 8)
        dump("amino8_sample", "sample", ["-guide", gf, "-tree", tf, "-preset", "lg", "-band", "4", "-profmaxstates", "400", "-seed", "7", "-profsamples", "6"])
        # 2. same family, posterior-profile mode (Forward+Backward+postProbProfile), no band
        dump("amino8_post", "post", ["-guide", gf, "-tree", tf, "-preset", "lg", "-noband", "-profminpost", ".01", "-profmaxstates", "400"], which="6,13,14")
        # 3. DNA, Jukes-Cantor with a 2-component gamma mixture, 6 leaves x ~60, band 6
        tf2, gf2 = simulate(tmp, balanced(6, 0.2), ["-preset", "jc", "-gamma", "2", "-rootlen", "60", "-seed", "5"], 6)
        dump("dna6_gamma2", "sample", ["-guide", gf2, "-tree", tf2, "-preset", "jc", "-gamma", "2", "-band", "6", "-profmaxstates", "500", "-seed", "11"])
    # 4. the reference's own test inputs: PF16593 (43 seqs x 32-36 aa) with its tree, band 10, root + a few nodes
    dump("pf16593_root", "sample", ["-guide", "data/PF16593.testspan.fa", "-tree", "data/PF16593.testspan.testnj.nh", "-model", "data/testamino.json",
                                    "-band", "10", "-profsamples", "10", "-profmaxstates", "300"], which="84")


if __name__ == "__main__":
    main()
