#!/usr/bin/env python3
"""Generates tests/golden/distmat_*.npz from the REAL reference: RateModel::distanceMatrix (model.cpp:506-655) and the
NJ / UPGMA trees built from it (tree.cpp:240-461), through oracle/_ref/bin/ref_dump_distmat.  Run in the build container."""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from historian_b200 import descs as D

BIN = os.path.join(ROOT, "oracle", "_ref", "bin")
OUT = os.path.join(ROOT, "tests", "golden")
REF = "/root/reference"


def leaf_alignment(tmp, preset_args, n, bl, rootlen, seed):
    import make_golden as G
    tf, gf = G.simulate(tmp, G.balanced(n, bl), preset_args + ["-rootlen", str(rootlen), "-seed", str(seed)], n)
    return gf   # leaf-only true alignment (the -guide file)


CASES = [
    ("distmat_gp120_lg", "lg", os.path.join(REF, "data", "gp120.guide.fa"), 100, None),
    ("distmat_lg12x150", "lg", None, 100, dict(preset_args=["-preset", "lg", "-insrate", "0.03", "-delrate", "0.03"], n=12, bl=0.2, rootlen=150, seed=6)),
    ("distmat_jc_gamma2_10x300", "jc:2:1", None, 100, dict(preset_args=["-preset", "jc", "-gamma", "2"], n=10, bl=0.15, rootlen=300, seed=2)),
    ("distmat_lg8x60_jconly", "lg", None, 0, dict(preset_args=["-preset", "lg"], n=8, bl=0.3, rootlen=60, seed=9)),
]


def main():
    for name, spec, fa, iters, sim in CASES:
        with tempfile.TemporaryDirectory() as tmp:
            if fa is None:
                fa = leaf_alignment(tmp, **sim)
            binf = os.path.join(tmp, "d.bin")
            subprocess.run([os.path.join(BIN, "ref_dump_distmat"), binf, spec, fa, str(iters)], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            d = D.read_dump(binf)
            d.pop("timing", None)   # wall-clock of this run: not part of the fixture
        d["preset"] = spec
        d["max_iterations"] = np.array([iters], np.int32)
        d = {k: (np.frombuffer(v.encode(), np.uint8) if isinstance(v, str) else v) for k, v in d.items()}
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
        print(name, d["dist"].shape, "rows", d["rows"].shape, "range %.4g .. %.4g" % (d["dist"][d["dist"] > 0].min(), d["dist"].max()),
              "%.1f kB" % (os.path.getsize(os.path.join(OUT, name + ".npz")) / 1e3))


if __name__ == "__main__":
    main()
