"""#=GF LP (TreeAlignFuncs::logLikelihood, sampler.cpp:372-449, recon.cpp:1138) against the reference binary's own
Stockholm output on BASELINE config 2 (PF16593).  Fixtures: tests/golden/pf16593/PF16593.recon[.ancseq].stk, written by
`oracle/_ref/bin/historian recon -seqs PF16593.fa -tree PF16593.nhx -output stockholm [-ancseq]` (tools/make_golden_r2.py)."""
import os

import numpy as np
import pytest

import helpers as H
import oracle_lib as O
from historian_b200 import descs as D, em
from historian_b200.model import RateModel
from historian_b200.simulator import Tree

FIX = [os.path.join(H.GOLDEN_DIR, "pf16593", f) for f in ("PF16593.recon.stk", "PF16593.recon.ancseq.stk")]


def load(path):
    names, rows, gf = em.read_stockholm_recon(path)
    tree = Tree.from_newick(gf["NH"])
    by_name = dict(zip(names, rows))
    cols = np.array([[ord(c) for c in by_name[n]] for n in tree.names], np.uint8).T.copy()   # [n_cols, N], node order
    # ancestors predicted with -ancseq are lower case (recon.cpp); the sum-product tokenises case-insensitively
    return tree, cols, float(gf["LP"])


@pytest.mark.parametrize("path", FIX, ids=["wildcard_ancestors", "ancseq"])
def test_gf_lp_with_the_cpu_oracle(path):
    tree, cols, want = load(path)
    model = RateModel.preset("lg")
    spm, topo = D.SumProdModel.from_rate_model(model), D.TreeTopology(tree.parent, tree.branch_length)
    cll = O.sumprod_columns(spm, topo, cols, flags=0)["col_loglike"]
    lp, *_ = em.tree_align_log_likelihood(model, tree, cols, cll)
    assert abs(lp - want) <= 1e-4 * abs(want) and abs(lp - want) < 5e-3, (lp, want)


@pytest.mark.gpu
@pytest.mark.parametrize("path", FIX, ids=["wildcard_ancestors", "ancseq"])
def test_gf_lp_on_device(ctx, path):
    tree, cols, want = load(path)
    model = RateModel.preset("lg")
    spm, topo = D.SumProdModel.from_rate_model(model), D.TreeTopology(tree.parent, tree.branch_length)
    sp = ctx.sumprod(spm, topo)
    try:
        cll = sp.columns(cols, flags=0)["col_loglike"]
    finally:
        sp.close()
    lp, lp_root, lp_gaps, lp_sub = em.tree_align_log_likelihood(model, tree, cols, cll)
    assert abs(lp - want) < 5e-3, (lp, want, lp_root, lp_gaps, lp_sub)
    ocl = O.sumprod_columns(spm, topo, cols, flags=0)["col_loglike"]
    assert np.allclose(cll, ocl, rtol=0, atol=1e-9)
