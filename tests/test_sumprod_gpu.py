"""Device sum-product / EM counts (historian_b200/csrc/hist_sumprod.cu) through the C ABI, against
(1) the CPU oracle on the same inputs and (2) the REAL reference's outputs in tests/golden/sumprod_*.npz.

Tolerances.  Device and oracle are both FP64 with the reference's operation order inside a message;
they differ in (a) device libm (exp/log/log1p, <= 2 ulp), (b) the summation order of the eigen-counts
over nodes and columns, (c) eigenSubCount multiplying by a precomputed reciprocal 1/(lambda_k-lambda_l).
So: messages and log scale factors 1e-12 relative, column log-likelihoods 1e-10 absolute, counts 1e-9
relative (to the largest count).  Against the reference's goldens the table log-sum-exp adds <= 4.6e-5 per
binary op when C > 1 (logsumexp.h:47): 1e-4 absolute on colLogLike / posteriors, 2e-4 relative on counts."""
import glob
import os

import numpy as np
import pytest

import helpers as H
import oracle_lib as O
from historian_b200 import capi, descs as D

pytestmark = pytest.mark.gpu
FIX = sorted(glob.glob(os.path.join(H.GOLDEN_DIR, "sumprod_*.npz")))
IDS = [os.path.basename(f)[8:-4] for f in FIX]


@pytest.fixture(scope="module")
def ctx():
    c = capi.Context(0)
    yield c
    c.close()


def load(path, use_branch_sub=True):
    d = H.load_golden(path)
    return d, D.SumProdModel.from_golden(d, use_branch_sub), D.TreeTopology(d["parent"], d["branch_length"])


def close(a, b, rel, abs_=0.0):
    a, b = np.asarray(a), np.asarray(b)
    inf = np.isneginf(b)
    assert H.same(np.isneginf(a), inf)
    return bool(np.all(np.abs(a[~inf] - b[~inf]) <= abs_ + rel * np.abs(b[~inf])))


def valid_masks(d):
    cols, parent = d["columns"], d["parent"]
    ungapped = ~np.isin(cols, [ord("-"), ord(".")])
    has_parent = parent >= 0
    par_ungapped = np.zeros_like(ungapped)
    par_ungapped[:, has_parent] = ungapped[:, parent[has_parent]]
    return ungapped, ~ungapped | par_ungapped


@pytest.mark.parametrize("path", FIX, ids=IDS)
def test_messages_match_oracle_and_reference(ctx, path):
    d, model, tree = load(path)
    sp = ctx.sumprod(model, tree)
    o = O.sumprod_columns(model, tree, d["columns"], debug=True)
    ungapped, e_valid = valid_masks(d)
    for k in range(len(d["columns"])):
        r = sp.debug_column(d["columns"][k])
        u = np.broadcast_to(ungapped[k][None, :, None], r["F"].shape)
        ev = np.broadcast_to(e_valid[k][None, :, None], r["E"].shape)
        u2 = np.broadcast_to(ungapped[k][None, :], r["logF"].shape)
        for ref in (o, d):
            assert close(r["F"][u], ref["F"][k][u], 1e-12, 1e-300)
            assert close(r["G"][u], ref["G"][k][u], 1e-12, 1e-300)
            assert close(r["E"][ev], ref["E"][k][ev], 1e-12, 1e-300)
            assert close(r["logF"], ref["logF"][k], 1e-12, 1e-12)
            assert close(r["logG"][u2], ref["logG"][k][u2], 1e-12, 1e-12)
        assert abs(r["col_loglike"] - o["col_loglike"][k]) <= 1e-10
    sp.close()


@pytest.mark.parametrize("path", FIX, ids=IDS)
def test_columns_counts_posteriors(ctx, path):
    d, model, tree = load(path)
    Cc, A, N = (int(v) for v in d["dims"])
    multi = Cc > 1
    sp = ctx.sumprod(model, tree)
    r = sp.columns(d["columns"], flags=capi.SP_COUNTS | capi.SP_MAP_STATES | capi.SP_NODE_POST)
    o = O.sumprod_columns(model, tree, d["columns"])
    # against the oracle
    assert close(r["col_loglike"], o["col_loglike"], 0, 1e-10)
    assert close(r["node_post"], o["node_post"], 0, 1e-9)
    assert H.same(r["map_state"], o["map_state"])
    scale = np.abs(o["counts"][7:]).max()
    assert np.all(np.abs(r["counts"] - o["counts"])[7:] <= 1e-9 * scale)
    assert abs(r["counts"][6] - o["counts"][6]) <= 1e-9 * len(d["columns"])
    assert np.all(r["counts"][:6] == 0)
    # against the reference itself
    assert close(r["col_loglike"], d["col_loglike"], 0, 1e-4 if multi else 1e-10)
    assert close(r["node_post"], d["node_post"], 0, 2e-4 if multi else 1e-9)
    assert H.same(r["map_state"], d["map_state"])
    _, root, ec = O.split_counts(r["counts"], Cc, A)
    ref_ec = d["cum_eigen_re"][-1] + 1j * d["cum_eigen_im"][-1]
    tol = 2e-4 if multi else 1e-9
    assert np.all(np.abs(root - d["cum_root_count"][-1]) <= tol * (1e-12 + np.abs(d["cum_root_count"][-1])))
    assert np.all(np.abs(ec - ref_ec) <= tol * (np.abs(ref_ec) + 1e-6 * np.abs(ref_ec).max()))
    # EigenModel::getSubCounts on device
    root2, sub = sp.transform_counts(r["counts"])
    assert np.all(np.abs(sub - d["sub_count_eigen"]) <= tol * (1e-6 + np.abs(d["sub_count_eigen"])))
    assert np.all(root2 == root)
    sp.close()


@pytest.mark.parametrize("path", FIX, ids=IDS)
def test_branch_tables_on_device(ctx, path):
    """S0: P(t) = V exp(Lambda t) V^-1 and eigenSubCount built on device from the eigen-system."""
    d, model, tree = load(path, use_branch_sub=False)
    sp = ctx.sumprod(model, tree)
    sub, esc = sp.branch_tables()
    osub, oesc = O.sumprod_branch_tables(model, tree)
    assert np.all(np.abs(sub - osub) <= 1e-13)
    assert np.all(np.abs(esc - oesc) <= 1e-12 * (1e-3 + np.abs(oesc)))
    ref_esc = d["branch_esc_re"] + 1j * d["branch_esc_im"]
    assert np.all(np.abs(esc - ref_esc) <= 1e-11 * (1e-3 + np.abs(ref_esc)))
    assert np.all(np.abs(sub - d["branch_sub"]) <= 1e-9)   # eigen route vs the reference's Pade expm (testmatexp: 10 digits)
    sp.close()


def test_weights_linearity_and_batch_independence(ctx):
    """Size-independent properties: counts are linear in the column weights, and the result does not depend
    on how columns are batched over blocks (many more columns than resident blocks)."""
    d, model, tree = load(FIX[-1])
    cols = d["columns"]
    sp = ctx.sumprod(model, tree)
    w = np.linspace(0.5, 2.0, len(cols))
    a = sp.columns(cols, w)["counts"]
    b = sum(w[k] * sp.columns(cols[k:k + 1])["counts"] for k in range(len(cols)))
    b[6] = sp.columns(cols)["counts"][6]
    assert np.allclose(a, b, rtol=1e-10, atol=1e-13)
    reps = 700
    big = np.tile(cols, (reps, 1))
    rb = sp.columns(big)
    one = sp.columns(cols)
    assert np.allclose(rb["counts"], reps * one["counts"], rtol=1e-10, atol=1e-12)
    assert np.array_equal(rb["col_loglike"], np.tile(one["col_loglike"], reps))
    sp.close()


def test_accumulates_into_caller_buffer_and_device_buffer(ctx):
    import torch
    d, model, tree = load(FIX[0])
    cols = d["columns"]
    sp = ctx.sumprod(model, tree)
    once = sp.columns(cols)["counts"]
    acc = np.zeros(sp.counts_size)
    sp.columns(cols, counts=acc)
    sp.columns(cols, counts=acc)
    assert np.allclose(acc, 2 * once, rtol=1e-13, atol=0)
    dev = torch.zeros(sp.counts_size, dtype=torch.float64, device="cuda:0")
    dcols = torch.from_numpy(np.ascontiguousarray(cols)).to("cuda:0")
    dll = torch.zeros(len(cols), dtype=torch.float64, device="cuda:0")
    sp.columns_device(dcols, len(cols), dev, dll)
    sp.columns_device(dcols, len(cols), dev, dll)
    ctx.synchronize()
    assert np.allclose(dev.cpu().numpy(), 2 * once, rtol=1e-13, atol=0)
    assert np.array_equal(dll.cpu().numpy(), sp.columns(cols)["col_loglike"])
    sp.close()


def test_multiple_roots_is_an_error_for_counts_but_not_for_likelihood(ctx):
    d, model, tree = load(FIX[0])
    N = tree.n_nodes
    col = np.full((1, N), ord("-"), np.uint8)
    leaves = [r for r in range(N) if r not in set(d["parent"].tolist())]
    col[0, leaves[0]] = d["alphabet"][0]
    col[0, leaves[-1]] = d["alphabet"][1]
    sp = ctx.sumprod(model, tree)
    r = sp.columns(col, flags=0)
    o = O.sumprod_columns(model, tree, col, flags=0)
    assert abs(r["col_loglike"][0] - o["col_loglike"][0]) <= 1e-10
    with pytest.raises(capi.HistError):
        sp.columns(col, flags=capi.SP_COUNTS)
    # an all-gap column contributes nothing
    empty = np.full((1, N), ord("-"), np.uint8)
    r = sp.columns(empty)
    assert np.all(r["counts"] == 0)
    sp.close()


@pytest.mark.parametrize("path", [f for f in FIX if "dna12" in f], ids=[i for f, i in zip(FIX, IDS) if "dna12" in f])
def test_dense_columns_node_major_path(ctx, path):
    """A batch large enough for the node-major kernels (dn_*: the dense columns of the batch processed together, level
    by level): counts and column likelihoods equal the oracle's, weights included, and equal the one-block-per-column
    kernel's (the same columns in a batch too small to be routed there)."""
    d, model, tree = load(path)
    rng = np.random.default_rng(1)
    cols = np.ascontiguousarray(np.tile(d["columns"], (12, 1)))
    rng.shuffle(cols)
    assert len(cols) >= 128
    w = rng.uniform(0.5, 2.0, len(cols))
    sp = ctx.sumprod(model, tree)
    launches0 = ctx.launch_count
    r = sp.columns(cols, weights=w, flags=capi.SP_COUNTS)
    n_launches = ctx.launch_count - launches0
    o = O.sumprod_columns(model, tree, cols, weights=w, flags=O.SP_COUNTS)
    assert close(r["col_loglike"], o["col_loglike"], 0, 1e-10)
    scale = np.abs(o["counts"][7:]).max()
    assert np.all(np.abs(r["counts"] - o["counts"])[7:] <= 1e-9 * scale)
    assert abs(r["counts"][6] - o["counts"][6]) <= 1e-9 * len(cols)
    # one launch per tree level and pass: the batch really went through the level-synchronous kernels
    depth = 0
    p = np.asarray(d["parent"])
    for v in range(len(p)):
        k, q = 0, v
        while p[q] >= 0:
            q, k = p[q], k + 1
        depth = max(depth, k + 1)
    assert n_launches >= 2 * depth, (n_launches, depth)
    # the same columns, a few at a time (column kernel only)
    acc = np.zeros_like(r["counts"])
    ll = []
    for k0 in range(0, len(cols), 40):
        part = sp.columns(cols[k0:k0 + 40], weights=w[k0:k0 + 40], flags=capi.SP_COUNTS)
        acc += part["counts"]
        ll.append(part["col_loglike"])
    assert close(r["col_loglike"], np.concatenate(ll), 0, 1e-10)
    assert np.all(np.abs(r["counts"] - acc)[7:] <= 1e-9 * scale)
    sp.close()
