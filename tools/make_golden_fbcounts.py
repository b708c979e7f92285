#!/usr/bin/env python3
"""Generates tests/golden/fbcounts_*.npz from the REAL reference: BackwardMatrix::getCounts
(forward.cpp:1183-1214) on two-leaf families, through oracle/_ref/bin/ref_dump_fbcounts
(oracle/ref_dump_fbcounts.cpp, linked against the unmodified reference objects).

Run in the build container (needs /root/reference and `make -C oracle ref`)."""
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

CASES = [
    # name, preset spec, generate args, branch lengths, band
    ("fbcounts_lg_band4", "lg", ["-preset", "lg", "-rootlen", "30", "-seed", "5", "-insrate", "0.06", "-delrate", "0.06"], (0.3, 0.5), 4),
    ("fbcounts_lg_noband", "lg", ["-preset", "lg", "-rootlen", "18", "-seed", "9", "-insrate", "0.08", "-delrate", "0.08"], (0.4, 0.2), -1),
    ("fbcounts_jc_gamma2_band3", "jc:2:1", ["-preset", "jc", "-gamma", "2", "-rootlen", "40", "-seed", "2", "-insrate", "0.05", "-delrate", "0.05"], (0.25, 0.25), 3),
]


TREE_CASES = [
    # name, leaves, branch length, preset, generate args, recon args: getCounts at the ROOT of a progressive reconstruction
    # (internal profiles with null states and multi-row alignment columns; transitions carry no counts)
    ("fbcounts_tree4_lg", 4, 0.25, "lg", ["-preset", "lg", "-rootlen", "24", "-seed", "4", "-insrate", "0.06", "-delrate", "0.06"],
     ["-preset", "lg", "-band", "5", "-profmaxstates", "300", "-seed", "3", "-profsamples", "8"]),
    ("fbcounts_tree6_jc_gamma2", 6, 0.2, "jc:2:1", ["-preset", "jc", "-gamma", "2", "-rootlen", "30", "-seed", "8"],
     ["-preset", "jc", "-gamma", "2", "-band", "4", "-profmaxstates", "300", "-seed", "5", "-profsamples", "6"]),
]
KEEP = ("x.", "y.", "hmm.", "env.", "fbc.")


def tree_cases(sfx="", propagate=False):
    """propagate: REF_DUMP_COUNTS=2, the strategy of `historian fit` on unaligned sequences (CountSubstEvents | CountIndelEvents):
    every transition of the internal profiles carries counts (forward.cpp:753-790) and getCounts reads them -- the fixtures
    fbcounts_prop_* of the part of row B3 that is not built yet."""
    import make_golden as G
    for name, n, bl, spec, gen, rec in TREE_CASES:
        name = name.replace("fbcounts_", "fbcounts_prop_") if propagate else name
        name += sfx
        with tempfile.TemporaryDirectory() as tmp:
            tf, gf = G.simulate(tmp, G.balanced(n, bl), gen, n)
            binf = os.path.join(tmp, "d.bin")
            root = 2 * n - 2
            subprocess.run([os.path.join(BIN, "ref_dump" + sfx), binf, str(root), "sample", "-guide", gf, "-tree", tf] + rec, check=True,
                           cwd=G.REF, env=dict(os.environ, REF_DUMP_COUNTS="2" if propagate else "1"), stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            d = D.read_dump(binf)
        pfx = "n%d." % root
        out = {"preset": spec, "root": np.array([root], np.int32)}
        for k, v in d.items():
            if k.startswith("tree.") or k in (pfx + "lp_end", pfx + "children") or (k.startswith(pfx) and k[len(pfx):].startswith(KEEP)):
                out[k] = v
        out = {k: (np.frombuffer(v.encode(), np.uint8) if isinstance(v, str) else v) for k, v in out.items()}
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
        print(name, "root", root, "x", out[pfx + "x.dims"], "y", out[pfx + "y.dims"], "indel", np.round(out[pfx + "fbc.indel"], 4),
              "%.1f kB" % (os.path.getsize(os.path.join(OUT, name + ".npz")) / 1e3))


def main(sfx=""):
    """sfx "": the reference as shipped (table log-sum-exp); "_exact": the same sources compiled with the reference's own
    -DLOG_SUM_EXP_SLOW switch (oracle/Makefile ref_exact) -> tests/golden/fbcounts_*_exact.npz, exact arithmetic."""
    tree_cases(sfx)
    tree_cases(sfx, propagate=True)
    for name, spec, gen, (tl, tr), band in CASES:
        name += sfx
        with tempfile.TemporaryDirectory() as tmp:
            tf = os.path.join(tmp, "tree.nh")
            open(tf, "w").write("(x:%g,y:%g);\n" % (tl, tr))
            out = subprocess.run([os.path.join(BIN, "historian"), "generate"] + gen + [tf, "-output", "fasta"],
                                 stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True).stdout.decode()
            seqs, cur = {}, None
            for line in out.splitlines():
                if line.startswith(">"):
                    cur = line[1:].strip()
                    seqs[cur] = ""
                elif cur:
                    seqs[cur] += line.strip()
            rows = {k: seqs[k] for k in ("x", "y")}
            keep = [i for i in range(len(rows["x"])) if any(v[i] not in "-." for v in rows.values())]
            gf = os.path.join(tmp, "pair.fa")
            with open(gf, "w") as f:
                for k, v in rows.items():
                    f.write(">%s\n%s\n" % (k, "".join(v[i] for i in keep)))
            binf = os.path.join(tmp, "d.bin")
            subprocess.run([os.path.join(BIN, "ref_dump_fbcounts" + sfx), binf, "preset:" + spec, gf, tf, str(band)], check=True,
                           stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            d = D.read_dump(binf)
        d = {k: (np.frombuffer(v.encode(), np.uint8) if isinstance(v, str) else v) for k, v in d.items()}
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
        print(name, "x", len(d["x_seq"]), "y", len(d["y_seq"]), "lp_end %.4f" % d["lp_end"][0], "indel", np.round(d["indel"], 4))




def elim_count_cases():
    """tests/golden/elimcounts_*_best.npz: every merge of the two tree families in `best` mode (one trace per profile) with the
    Count* strategy bits: inputs of makeProfile (child profiles with the indel counts their transitions carry, PairHMM,
    envelope, the best trace = the cell set) and its output (the parent's transitions and THEIR indel counts,
    forward.cpp:753-790) -- the pin of oracle_eliminate_indel_counts."""
    import make_golden as G
    keep = ("x.", "y.", "hmm.", "env.", "best_trace", "parent.dims", "parent.trans_src", "parent.trans_dest", "parent.trans_lp", "parent.trans_indel",
            "fbc.waits", "children", "lp_end")
    for name, n, bl, spec, gen, rec in TREE_CASES:
        with tempfile.TemporaryDirectory() as tmp:
            tf, gf = G.simulate(tmp, G.balanced(n, bl), gen, n)
            binf = os.path.join(tmp, "d.bin")
            subprocess.run([os.path.join(BIN, "ref_dump"), binf, "all", "best", "-guide", gf, "-tree", tf] + rec, check=True, cwd=G.REF,
                           env=dict(os.environ, REF_DUMP_COUNTS="2"), stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            d = D.read_dump(binf)
        out = {"preset": spec, "dumped_nodes": d["dumped_nodes"], "strategy": d["strategy"]}
        for k, v in d.items():
            if k.startswith("n") and "." in k and k.split(".", 1)[1].startswith(keep):
                if k.endswith(("trans_root_count", "trans_sub_count", "align_col", "lp_absorb")) and not k.endswith("lp_absorb"):
                    continue
                out[k] = v
        out = {k: (np.frombuffer(v.encode(), np.uint8) if isinstance(v, str) else v) for k, v in out.items()}
        fn = os.path.join(OUT, name.replace("fbcounts_", "elimcounts_") + "_best.npz")
        np.savez_compressed(fn, **out)
        print(os.path.basename(fn), "nodes", list(d["dumped_nodes"]), "%.1f kB" % (os.path.getsize(fn) / 1e3))


if __name__ == "__main__":
    main()
    main("_exact")
    elim_count_cases()
