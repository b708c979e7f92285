"""Distance-matrix stage: RateModel::distanceMatrix (model.cpp:506-655).

Fixtures tests/golden/distmat_*.npz come from the REAL reference (oracle/ref_dump_distmat.cpp through
tools/make_golden_distmat.py): its own data/gp120.guide.fa under LG, simulated LG and JC+gamma families, and a
Jukes-Cantor-only run (maxIterations = 0).  The reference exponentiates with GSL's scaling-and-squaring routine, the
restatement and the device with the model's eigen-decomposition (~1e-13 apart), so a golden-section comparison can flip
only on a tie at that precision: results must agree to 1e-6 relative for at least 95 % of the pairs and to the search's
own convergence tolerance (1 %) for all of them."""
import glob
import os

import numpy as np
import pytest

import helpers as H
import oracle_lib as O
from historian_b200 import descs as D, guide as G
from historian_b200.model import RateModel

FIX = sorted(glob.glob(os.path.join(H.GOLDEN_DIR, "distmat_*.npz")))
IDS = [os.path.basename(f)[8:-4] for f in FIX]


def load(path):
    d = H.load_golden(path)
    f = bytes(d["preset"]).decode().split(":")
    m = RateModel.preset(f[0])
    if len(f) == 3:
        m = m.gamma(int(f[1]), float(f[2]))
    return d, m, int(d["max_iterations"][0])


def check(got, want, iters):
    assert got.shape == want.shape and np.array_equal(got, got.T) and np.all(np.diag(got) == 0)
    iu = np.triu_indices(len(want), 1)
    g, w = got[iu], want[iu]
    rel = np.abs(g - w) / np.abs(w)
    if iters <= 0:
        assert np.all(rel <= 1e-9), rel.max()
        return
    assert np.all(rel <= 0.0101), (rel.max(), g[np.argmax(rel)], w[np.argmax(rel)])
    assert np.mean(rel <= 1e-6) >= 0.95, (np.mean(rel <= 1e-6), np.sort(rel)[-5:])


@pytest.mark.parametrize("path", FIX, ids=IDS)
def test_oracle_restatement_matches_the_reference(path):
    d, m, iters = load(path)
    rate = G.expected_substitution_rate(m)
    assert abs(rate - d["expected_sub_rate"][0]) <= 1e-12
    got = O.distance_matrix(D.SumProdModel.from_rate_model(m), rate, d["rows"], iters)
    check(got, d["dist"], iters)


@pytest.mark.gpu
@pytest.mark.parametrize("path", FIX, ids=IDS)
def test_device_matches_the_reference_and_the_oracle(ctx, path):
    d, m, iters = load(path)
    got = G.distance_matrix(ctx, m, d["rows"], iters)
    check(got, d["dist"], iters)
    want = O.distance_matrix(D.SumProdModel.from_rate_model(m), G.expected_substitution_rate(m), d["rows"], iters)
    check(got, want, iters)
    assert np.array_equal(got, G.distance_matrix(ctx, m, d["rows"], iters))   # reproducible


@pytest.mark.gpu
def test_edge_cases(ctx):
    """Rows that share no column (no pair counts at all: the reference's scan never finds a bracket and returns a bound),
    identical rows, wildcards and lower / upper case, a single row."""
    m = RateModel.preset("lg")
    spm, rate = D.SumProdModel.from_rate_model(m), G.expected_substitution_rate(m)
    rows = np.array([list(b"ARND--------"), list(b"--------cqeg"), list(b"ARNDCQEGHILK"), list(b"arndcqeg*ilk"), list(b"VVVVVVVVVVVV")], np.uint8)
    got = ctx.distance_matrix(spm, rate, rows)
    want = O.distance_matrix(spm, rate, rows)
    assert np.allclose(got, want, rtol=1e-6, atol=0), (got, want)
    assert got[0, 1] == 10.0 and got[0, 1] == want[0, 1]   # no counts: the scan finds no bracket, tUpper comes back
    assert got[2, 3] < 1e-3                      # identical up to case and a wildcard
    assert np.array_equal(ctx.distance_matrix(spm, rate, rows[:1]), np.zeros((1, 1)))


@pytest.mark.gpu
@pytest.mark.parametrize("path", FIX, ids=IDS)
@pytest.mark.parametrize("method", ["upgma", "nj"])
def test_trees_equal_the_reference(ctx, path, method):
    """Tree::buildByUPGMA / buildByNeighborJoining (tree.cpp:240-461) from the reference's own distance matrix: the Newick text
    (joins, child order, 6-digit branch lengths) must be the reference's, character for character."""
    d, m, iters = load(path)
    names = bytes(d["names"]).decode().split("\n")[:-1]
    got = G.build_tree(ctx, names, d["dist"], method)
    assert got == bytes(d[method]).decode()


@pytest.mark.parametrize("path", FIX, ids=IDS)
@pytest.mark.parametrize("method", ["upgma", "nj"])
def test_oracle_trees_equal_the_reference(path, method):
    d, m, iters = load(path)
    names = bytes(d["names"]).decode().split("\n")[:-1]
    child, blen = O.build_tree(d["dist"], 0 if method == "upgma" else 1)
    assert G.newick(names, child, blen) == bytes(d[method]).decode()


@pytest.mark.gpu
def test_device_tree_is_bit_identical_to_the_oracle_on_a_large_matrix(ctx):
    """300 rows (more than any fixture), distances with many exact ties (multiples of 1/8) so that the first-minimum rule decides
    joins: children and branch lengths must be equal bit for bit, for both methods."""
    rng = np.random.default_rng(3)
    n = 300
    a = rng.integers(1, 40, (n, n)) / 8.0
    dist = np.triu(a, 1) + np.triu(a, 1).T
    for method in (0, 1):
        child, blen = ctx.build_tree(dist, method)
        wc, wb = O.build_tree(dist, method)
        assert np.array_equal(child, wc), method
        assert np.array_equal(blen, wb), method
