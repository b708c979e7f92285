"""Pins the sum-product oracle (oracle/oracle_sumprod.cpp) against the REAL reference: every fixture
tests/golden/sumprod_*.npz was produced by the unmodified AlignColSumProduct / EigenModel
(oracle/_ref/bin/ref_dump_sumprod, tools/make_golden_sumprod.py).  Runs on CPU.

Tolerances: messages, scale factors, posteriors and counts are FP64 on both sides, computed with the
same operations in the same order except (a) the reference's table log-sum-exp (<= 4.6e-5 per binary
op, logsumexp.h:47) inside colLogLike / logNodePostProb when C > 1, and (b) complex multiply
association.  So: E/F/G/log scale factors 1e-12 relative; col_loglike / node_post 1e-4 absolute when
C > 1 else 1e-12; counts 1e-4 relative when C > 1 (they are normalised by exp(colLogLike)) else 1e-10."""
import glob
import os

import numpy as np
import pytest

import helpers as H
import oracle_lib as O
from historian_b200 import descs as D

FIX = sorted(glob.glob(os.path.join(H.GOLDEN_DIR, "sumprod_*.npz")))


def load(path):
    d = H.load_golden(path)
    return d, D.SumProdModel.from_golden(d), D.TreeTopology(d["parent"], d["branch_length"])


def close(a, b, rel, abs_=0.0):
    a, b = np.asarray(a), np.asarray(b)
    inf = np.isneginf(b)
    assert H.same(np.isneginf(a), inf)
    return bool(np.all(np.abs(a[~inf] - b[~inf]) <= abs_ + rel * np.abs(b[~inf])))


@pytest.mark.parametrize("path", FIX, ids=[os.path.basename(f)[8:-4] for f in FIX])
def test_columns_match_reference(path):
    d, model, tree = load(path)
    Cc, A, N = (int(v) for v in d["dims"])
    r = O.sumprod_columns(model, tree, d["columns"], debug=True)
    multi = Cc > 1
    # messages of the ungapped nodes (gapped rows keep stale values in the reference)
    ungapped = ~np.isin(d["columns"], [ord("-"), ord(".")])                       # [L, N]
    m4 = np.broadcast_to(ungapped[:, None, :, None], r["F"].shape)
    for k in ("F", "G"):
        assert close(r[k][m4], d[k][m4], 1e-12, 1e-300), k
    # E of a column root is never written by fillUp: compare non-root ungapped rows and gapped rows (= 1)
    parent = d["parent"]
    has_parent = np.array([p >= 0 for p in parent])
    par_ungapped = np.zeros_like(ungapped)
    par_ungapped[:, has_parent] = ungapped[:, parent[has_parent]]
    e_valid = ~ungapped | par_ungapped
    me = np.broadcast_to(e_valid[:, None, :, None], r["E"].shape)
    assert close(r["E"][me], d["E"][me], 1e-12, 1e-300)
    m3 = np.broadcast_to(ungapped[:, None, :], r["logF"].shape)
    assert close(r["logF"], d["logF"], 1e-12, 1e-12)
    assert close(r["logG"][m3], d["logG"][m3], 1e-12, 1e-12)
    assert close(r["col_loglike"], d["col_loglike"], 0, 1e-4 if multi else 1e-11)
    assert close(r["node_post"], d["node_post"], 0, 2e-4 if multi else 1e-10)
    assert H.same(r["map_state"], d["map_state"])
    _, root, ec = O.split_counts(r["counts"], Cc, A)
    ref_ec = d["cum_eigen_re"][-1] + 1j * d["cum_eigen_im"][-1]
    tol = 2e-4 if multi else 1e-10
    assert np.all(np.abs(root - d["cum_root_count"][-1]) <= tol * (1e-12 + np.abs(d["cum_root_count"][-1])))
    assert np.all(np.abs(ec - ref_ec) <= tol * (np.abs(ref_ec) + 1e-9 * np.abs(ref_ec).max()))
    assert abs(r["counts"][6] - float(d["lp_total"][0])) <= (1e-4 if multi else 1e-10) * len(d["col_loglike"])
    # EigenModel::getSubCounts on the reference's own eigen-counts, and on ours
    full = np.zeros(D.counts_size(Cc, A))
    full[7:7 + Cc * A] = d["cum_root_count"][-1].ravel()
    full[7 + Cc * A:] = np.stack([ref_ec.real, ref_ec.imag], -1).ravel()
    _, sub = O.sumprod_transform_counts(model, full)
    assert np.all(np.abs(sub - d["sub_count_eigen"]) <= 1e-10 * (1e-6 + np.abs(d["sub_count_eigen"])))
    _, sub2 = O.sumprod_transform_counts(model, r["counts"])
    assert np.all(np.abs(sub2 - d["sub_count_eigen"]) <= tol * (1e-6 + np.abs(d["sub_count_eigen"])))


@pytest.mark.parametrize("path", FIX, ids=[os.path.basename(f)[8:-4] for f in FIX])
def test_branch_tables_from_eigen(path):
    """eigenSubCount is exact w.r.t. the reference; exp(Rt) via the eigen route agrees with the reference's
    Pade expm to the 10 digits its own testmatexp golden asks for (Makefile:221-223)."""
    d, model, tree = load(path)
    sub, esc = O.sumprod_branch_tables(model, tree)
    ref_esc = d["branch_esc_re"] + 1j * d["branch_esc_im"]
    assert np.all(np.abs(esc - ref_esc) <= 1e-12 * (1e-3 + np.abs(ref_esc)))
    assert np.all(np.abs(sub - d["branch_sub"]) <= 1e-9)


def test_weights_and_linearity():
    d, model, tree = load(FIX[-1])
    cols = d["columns"]
    w = np.linspace(0.5, 2.0, len(cols))
    a = O.sumprod_columns(model, tree, cols, w, flags=O.SP_COUNTS)["counts"]
    b = sum(w[k] * O.sumprod_columns(model, tree, cols[k:k + 1], flags=O.SP_COUNTS)["counts"] for k in range(len(cols)))
    b[6] = O.sumprod_columns(model, tree, cols, flags=O.SP_COUNTS)["counts"][6]  # lp is not weighted (model.cpp:909)
    assert np.allclose(a, b, rtol=1e-10, atol=1e-14)
