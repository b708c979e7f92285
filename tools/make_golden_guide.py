#!/usr/bin/env python3
"""Generates tests/golden/guide_*.npz from the REAL reference: AlignGraph (span.cpp:35-158: random or all-pairs graph of
QuickAlignMatrix alignments, spanning tree from sequence 0, alignPathMerge) through oracle/_ref/bin/ref_dump_guide
(oracle/ref_dump_guide.cpp, linked against the unmodified reference objects).  The random graph depends on the standard
library's uniform_int_distribution (the reason the reference skips its own `make testspan`): these fixtures are what the
reference compiled HERE (g++ / libstdc++) produces.  Run in the build container (needs `make -C oracle ref`)."""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from make_golden_quickalign import simulate, BIN, OUT, REF

CASES = [
    # name, preset, fasta (None: simulate), time, graph, seed, envelope mode
    ("guide_pf16593_rnd", "lg", os.path.join(REF, "data", "PF16593.fa"), 1.0, "rnd", 5489, "sparse:64:6:-1:%d" % (1 << 34), None),
    ("guide_pf16593_all", "lg", os.path.join(REF, "data", "PF16593.fa"), 1.0, "all", 5489, "full", None),
    ("guide_lg24x200_rnd_seed7", "lg", None, 1.0, "rnd", 7, "full", dict(n=24, bl=0.08, rootlen=200, seed=11, extra=["-insrate", "0.04", "-delrate", "0.04"])),
    ("guide_lg12x400_kmatchn", "lg", None, 1.0, "rnd", 5489, "sparse:16:5:2:0", dict(n=12, bl=0.04, rootlen=400, seed=12, extra=["-insrate", "0.03", "-delrate", "0.03"])),
    ("guide_jc10x300_kmatchmb", "jc", None, 1.0, "all", 5489, "sparse:32:6:-1:200000", dict(n=10, bl=0.05, rootlen=300, seed=13, extra=["-insrate", "0.03", "-delrate", "0.03"])),
]


def read_fasta(text):
    names, seqs = [], []
    for line in text.splitlines():
        if line.startswith(">"):
            names.append(line[1:].split()[0])
            seqs.append("")
        elif names:
            seqs[-1] += line.strip()
    return names, seqs


def main():
    for name, preset, fa, time, graph, seed, mode, sim in CASES:
        with tempfile.TemporaryDirectory() as tmp:
            if fa is None:
                fa = simulate(tmp, preset, **sim)
            r = subprocess.run([os.path.join(BIN, "ref_dump_guide"), preset, fa, str(time), graph, str(seed), mode], check=True,
                               stdout=subprocess.PIPE, stderr=subprocess.PIPE)
            _, seqs = read_fasta(open(fa).read())
        names, gapped = read_fasta(r.stdout.decode())
        d = {"graph": np.frombuffer(graph.encode(), np.uint8), "seed": np.array([seed]), "mode": np.frombuffer(mode.encode(), np.uint8)}
        edges, lps = [], []
        for line in r.stderr.decode().splitlines():
            w = line.split()
            if w[0] == "alphabet":
                d["alphabet"] = np.frombuffer(w[1].encode(), np.uint8)
            elif w[0] == "scores":
                d["scores"] = np.array([float(v) for v in w[1:]])
            elif w[0] == "submat":
                d["submat"] = np.array([float(v) for v in w[1:]])
            elif w[0] == "edge":
                edges.append((int(w[1]), int(w[2])))
                lps.append(float(w[3]))
        d["edges"], d["edge_lp"] = np.array(edges, np.int32), np.array(lps)
        d["seqs"] = np.frombuffer("\n".join(seqs).encode(), np.uint8)
        d["gapped"] = np.frombuffer("\n".join(gapped).encode(), np.uint8)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
        print(name, "seqs", len(seqs), "edges", len(edges), "columns", len(gapped[0]), "%.1f kB" % (os.path.getsize(os.path.join(OUT, name + ".npz")) / 1e3))


if __name__ == "__main__":
    main()
