#!/usr/bin/env python3
"""Generates tests/golden/quickalign_*.npz from the REAL reference: QuickAlignMatrix (quickalign.cpp:7-214) inside full
and sparse (k-mer seeded) DiagonalEnvelopes, through oracle/_ref/bin/ref_dump_quickalign (oracle/ref_dump_quickalign.cpp,
linked against the unmodified reference objects).  Run in the build container (needs `make -C oracle ref`)."""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from historian_b200 import descs as D

BIN = os.path.join(ROOT, "oracle", "_ref", "bin")
OUT = os.path.join(ROOT, "tests", "golden")
REF = "/root/reference"


def simulate(tmp, preset, n, bl, rootlen, seed, extra=()):
    """n leaf sequences of a star-ish random family from the reference's own simulator."""
    def rec(lo, hi):
        if hi - lo == 1:
            return "s%d:%g" % (lo, bl)
        mid = (lo + hi) // 2
        return "(%s,%s):%g" % (rec(lo, mid), rec(mid, hi), bl)
    tf = os.path.join(tmp, "t.nh")
    open(tf, "w").write(rec(0, n).rsplit(":", 1)[0] + ";\n")
    out = subprocess.run([os.path.join(BIN, "historian"), "generate", "-preset", preset, "-rootlen", str(rootlen), "-seed", str(seed)] + list(extra)
                         + [tf, "-output", "fasta"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True).stdout.decode()
    seqs, cur = {}, None
    for line in out.splitlines():
        if line.startswith(">"):
            cur = line[1:].strip()
            seqs[cur] = ""
        elif cur:
            seqs[cur] += line.strip()
    fa = os.path.join(tmp, "leaves.fa")
    with open(fa, "w") as f:
        for k in range(n):
            f.write(">s%d\n%s\n" % (k, seqs["s%d" % k].replace("-", "").replace(".", "")))
    return fa


CASES = [
    # name, preset, fasta (None: simulate with the args that follow), time, envelope mode
    ("quickalign_pf16593_pair_full", "lg", os.path.join(REF, "data", "PF16593.pair.fa"), 1.0, "full", None),
    ("quickalign_lg6x120_full", "lg", None, 1.0, "full", dict(n=6, bl=0.15, rootlen=120, seed=3, extra=["-insrate", "0.05", "-delrate", "0.05"])),
    ("quickalign_lg8x300_sparse", "lg", None, 1.0, "sparse:16:5:2", dict(n=8, bl=0.05, rootlen=300, seed=5, extra=["-insrate", "0.03", "-delrate", "0.03"])),
    ("quickalign_jc5x200_sparse", "jc", None, 0.5, "sparse:8:6:1", dict(n=5, bl=0.1, rootlen=200, seed=7, extra=["-insrate", "0.04", "-delrate", "0.04"])),
    # the memory-bounded threshold (-kmatchmb): kmerThreshold -1, the largest set of seeded bands that fits 400 kB at 40 B per cell
    ("quickalign_lg6x250_kmatchmb", "lg", None, 1.0, "sparse:12:5:-1:400000", dict(n=6, bl=0.05, rootlen=250, seed=21, extra=["-insrate", "0.03", "-delrate", "0.03"])),
]


def main():
    for name, preset, fa, time, mode, sim in CASES:
        with tempfile.TemporaryDirectory() as tmp:
            if fa is None:
                fa = simulate(tmp, preset, **sim)
            binf = os.path.join(tmp, "d.bin")
            subprocess.run([os.path.join(BIN, "ref_dump_quickalign"), binf, preset, fa, str(time), mode], check=True,
                           stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            d = D.read_dump(binf)
            d.pop("timing", None)   # wall-clock of this run: not part of the fixture
        d = {k: (np.frombuffer(v.encode(), np.uint8) if isinstance(v, str) else v) for k, v in d.items()}
        for k in list(d):   # tokens fit int8
            if k.endswith("_tok"):
                d[k] = d[k].astype(np.int8)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
        n = int(d["n_pairs"][0])
        nd = [len(d["p%d.diagonals" % k]) for k in range(n)]
        print(name, "pairs", n, "diagonals", min(nd), "-", max(nd), "finite", sum(np.isfinite(d["p%d.result" % k][0]) for k in range(n)),
              "%.1f kB" % (os.path.getsize(os.path.join(OUT, name + ".npz")) / 1e3))


if __name__ == "__main__":
    main()
