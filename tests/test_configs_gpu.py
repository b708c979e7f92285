"""BASELINE configs 1 and 2 on their own inputs, whole tree, against the reference BINARY (oracle/_ref/bin/historian,
run at test time on the same box) -- both drivers (per-merge Python driver and the native level-batched one):

  config 1  data/gp120.fa (10 HIV/SIV GP120 sequences, ~500 aa) with its guide alignment and tree
  config 2  data/PF16593.fa (43 sequences x 32-36 aa) with data/PF16593.nhx and the guide the reference itself
            computes (saved once with -saveguide by tools/make_golden_r2.py; the guide aligner is outside the hot path)

in deterministic mode (-profmaxstates 1) and in the DEFAULT sampling mode (-profsamples 10, seed 5489, one mt19937
stream in node order), with -profmaxstates fixed (the reference derives it from host RAM otherwise).  The reconstructed
alignment must be identical row by row.  The sampled runs walk hundreds of stochastic traceback steps through matrices
that differ from the reference's by its table-log-sum-exp error, so a draw may land on the other side of a boundary;
should that happen the test reports how many rows differ instead of hiding it (it does not on these inputs).
"""
import os
import subprocess

import numpy as np
import pytest

import helpers as H
from historian_b200 import model as M, recon as R, simulator as S
from test_recon_gpu import _read_fasta, REF_BIN

pytestmark = pytest.mark.gpu
G = H.GOLDEN_DIR
CONFIGS = {
    "gp120": (os.path.join(G, "gp120", "gp120.guide.fa"), os.path.join(G, "gp120", "gp120.tree.nh")),
    "pf16593": (os.path.join(G, "pf16593", "PF16593.refguide.fa"), os.path.join(G, "pf16593", "PF16593.nhx")),
}


def _load(name):
    gf, tf = CONFIGS[name]
    tree = S.Tree.from_newick(open(tf).read())
    rows = dict(_read_fasta(open(gf).read()))
    node_of = {nm: v for v, nm in enumerate(tree.names) if tree.is_leaf(v)}
    assert set(rows) == set(node_of), "guide rows and tree leaves differ"
    seqs = {node_of[nm]: r.replace("-", "").replace(".", "") for nm, r in rows.items()}
    grows = {node_of[nm]: np.array([c not in "-." for c in r]) for nm, r in rows.items()}
    return tree, seqs, grows, gf, tf


def _reference(gf, tf, extra):
    args = [REF_BIN, "recon", "-guide", gf, "-tree", tf, "-output", "fasta", "-norefine"] + extra
    out = subprocess.run(args, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True).stdout.decode()
    return _read_fasta(out)


def _check(ref_rows, ours, tree, what):
    assert len(ref_rows) == tree.n_nodes
    diff = [k for k, (name, row) in enumerate(ref_rows) if not (ours[k] == row or ours[k].lower() == row.lower())]
    assert not diff, "%s: %d of %d rows differ from the reference's (first: node %d)\n%s\n%s" % (
        what, len(diff), tree.n_nodes, diff[0], ours[diff[0]], ref_rows[diff[0]][1])


@pytest.mark.parametrize("name,max_states,samples", [("gp120", 1, 10), ("gp120", 1500, 10), ("pf16593", 1, 10), ("pf16593", 300, 10), ("pf16593", 0, 3)])
def test_whole_tree_matches_reference_binary(ctx, name, max_states, samples):
    if not os.path.exists(REF_BIN):
        pytest.skip("oracle/_ref/bin/historian not built")
    tree, seqs, grows, gf, tf = _load(name)
    ref_rows = _reference(gf, tf, ["-profmaxstates", str(max_states), "-profsamples", str(samples)])
    lg = M.RateModel.preset("lg")   # the reference's default amino-acid model (recon.cpp:671-680)
    nat = R.NativeReconstructor(ctx, lg, band=20, profile_samples=samples, max_profile_states=max_states)
    bits, lp_nat, rec = nat.reconstruct(tree, seqs, grows)
    _check(ref_rows, R.gapped_rows_dense(bits, tree, seqs, wildcard="x"), tree, name + " native")
    rec.close()
    if name == "pf16593" or max_states == 1:   # the per-merge driver too (slow on the large sampled profiles of gp120)
        py = R.Reconstructor(ctx, lg, band=20, profile_samples=samples, max_profile_states=max_states)
        path, lp_py, _ = py.reconstruct(tree, seqs, grows)
        _check(ref_rows, R.gapped_rows(path, tree, seqs, wildcard="x"), tree, name + " python")
        assert lp_py == lp_nat
