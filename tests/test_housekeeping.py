"""Small pins the round-1 review asked for.

  * gsl_cdf_gamma_Pinv of the oracle's GSL shim (no reference golden uses -gamma) against scipy.stats.gamma.ppf;
  * the numpy simulator (historian_b200/simulator.py, the generator of the benchmark families) against the reference's
    own `historian generate` on a 64-leaf tree: same model, different random streams, so the comparison is statistical
    (alignment length, indel-run lengths, sibling identity)."""
import ctypes
import os
import subprocess
import tempfile

import numpy as np
import pytest

import helpers as H
import oracle_lib as O
from historian_b200 import model as M, simulator as S

REF_BIN = os.path.join(os.path.dirname(H.GOLDEN_DIR), "..", "oracle", "_ref", "bin", "historian")


def test_gamma_pinv_shim_matches_scipy():
    from scipy.stats import gamma
    f = O.lib().oracle_gamma_pinv
    f.restype = ctypes.c_double
    f.argtypes = [ctypes.c_double] * 3
    for shape in (0.1, 0.5, 1.0, 2.0, 7.5):
        for p in (0.001, 0.1, 1 / 3, 0.5, 2 / 3, 0.9, 0.999):
            want = gamma.ppf(p, a=shape, scale=1.0 / shape)
            got = f(p, shape, 1.0 / shape)
            assert abs(got - want) <= 1e-9 * max(1.0, abs(want)), (shape, p, got, want)


def _stats(rows):
    """rows: uint8 [n_rows][cols] of a leaf alignment ('-' = gap)."""
    gap = rows == ord("-")
    runs = []
    for r in gap:
        d = np.diff(np.concatenate([[0], r.astype(np.int8), [0]]))
        runs.extend((np.nonzero(d == -1)[0] - np.nonzero(d == 1)[0]).tolist())
    ident = []
    for a, b in zip(rows[0::2], rows[1::2]):
        both = (a != ord("-")) & (b != ord("-"))
        ident.append(float((a[both] == b[both]).mean()))
    return rows.shape[1], float(np.mean(runs)) if runs else 0.0, float(np.mean(gap)), float(np.mean(ident))


def test_numpy_simulator_matches_reference_generate():
    if not os.path.exists(REF_BIN):
        pytest.skip("oracle/_ref/bin/historian not built")
    n_leaves, root_len, reps = 64, 300, 4
    tree = S.Tree.balanced(n_leaves, 0.1)
    lg = M.RateModel.preset("lg")
    lg.insRate = lg.delRate = 0.03
    ours, theirs = [], []
    with tempfile.TemporaryDirectory() as tmp:
        tf = os.path.join(tmp, "tree.nh")
        open(tf, "w").write(tree.newick())
        for seed in range(1, reps + 1):
            gapped, _ = S.simulate_family(lg, tree, root_len, seed)
            ours.append(_stats(S.leaf_guide(tree, gapped)[1]))
            out = subprocess.run([REF_BIN, "generate", "-preset", "lg", "-insrate", "0.03", "-delrate", "0.03", "-rootlen", str(root_len),
                                  "-seed", str(seed), tf, "-output", "fasta"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True).stdout.decode()
            rows, name = {}, None
            for line in out.splitlines():
                if line.startswith(">"):
                    name = line[1:].strip()
                    rows[name] = ""
                elif name is not None:
                    rows[name] += line.strip()
            leaves = np.array([list(rows[tree.names[v]].encode()) for v in tree.leaves()], np.uint8)
            leaves[leaves == ord(".")] = ord("-")
            leaves = leaves[:, (leaves != ord("-")).any(axis=0)]
            theirs.append(_stats(np.char.lower(leaves.view("S1")).view(np.uint8).reshape(leaves.shape)))
    a, b = np.mean(ours, axis=0), np.mean(theirs, axis=0)
    # columns, mean gap-run length, gap fraction, sibling identity: means over the replicates within 10 % (5 % for identity)
    for k, tol in enumerate((0.10, 0.10, 0.15, 0.05)):
        assert abs(a[k] - b[k]) <= tol * abs(b[k]), (k, a, b)
