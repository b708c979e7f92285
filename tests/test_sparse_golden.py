"""Parity on BASELINE config 1 (data/gp120.fa with its guide and tree: root + two internal nodes) and on a full-size
internal-node merge (a depth-3 node of a 16-leaf LG family: ~1000-state DAG children with in-degree up to 4,
~300 in-envelope cells per row, so the anti-diagonals run past 32 and 96 slots and the general fill kernel is used).

The reference's matrices of these nodes are too large to commit densely; the fixtures hold every stride-th in-envelope
cell (row-major order) with its five values, and the number of in-envelope cells of every row (tools/make_golden_r2.py).

  CPU  (not gpu): the oracle restatement reproduces those cells, the likelihoods and the traces bit for bit.
  GPU           : device tables / fills / traces against the reference's values (tolerances of test_fill_gpu.py) and
                  against the exact-log-sum-exp oracle over the WHOLE matrix.
"""
import numpy as np
import pytest

import helpers as H
import oracle_lib as O
from historian_b200 import descs as D

CASES = H.golden_cases(sparse=True)
REL_GOLD, ABS_GOLD = 1e-4, 5e-3
REL_EXACT, ABS_EXACT = 2e-6, 2e-4


@pytest.fixture(scope="module")
def goldens():
    return {f: H.load_golden(f) for f in H.golden_files(sparse=True)}


def test_have_sparse_fixtures():
    assert len(CASES) >= 4


def _sparse(d, pfx, which):
    xy = d[pfx + which + "_xy"]
    return xy[:, 0], xy[:, 1], d[pfx + which + "_vals"]


@pytest.mark.parametrize("cid,path,node", CASES, ids=[c[0] for c in CASES])
def test_oracle_reproduces_reference_sparse(goldens, cid, path, node):
    d = goldens[path]
    pfx = "n%d." % node
    x, y, hmm, env = H.node_inputs(d, node)
    t = O.pair_tables(x, y, hmm, env)
    for k in ("insx", "insy", "rootsubx", "rootsuby", "x_near_start", "y_near_end"):
        assert H.same(t[k], d[pfx + k]), k
    assert H.same(t["in_envelope"].sum(axis=1).astype(np.int32), d[pfx + "fwd_row_count"])
    fwd, lp_end = O.forward(x, y, hmm, env)
    i, j, v = _sparse(d, pfx, "fwd")
    assert H.same(fwd[i, j], v) and lp_end == d[pfx + "lp_end"][0]
    bwd, lp_start = O.backward(x, y, hmm, env)
    i, j, v = _sparse(d, pfx, "bwd")
    assert H.same(bwd[i, j], v) and lp_start == d[pfx + "lp_start"][0]
    fwd = np.ascontiguousarray(fwd)
    bt = O.trace_best(x, y, hmm, env, fwd, (x.n_states - 1, y.n_states - 1, D.EEE))
    assert H.same(bt, d[pfx + "best_trace"])
    traces, st = O.trace_sample(x, y, hmm, env, fwd, d[pfx + "rng_before"], 3)
    for k in range(3):
        assert H.same(traces[k], d[pfx + "sample_trace%d" % k])
    assert H.same(st, d[pfx + "rng_after"])


@pytest.mark.gpu
@pytest.mark.parametrize("cid,path,node", CASES, ids=[c[0] for c in CASES])
def test_device_matches_reference_sparse(ctx, goldens, cid, path, node):
    d = goldens[path]
    pfx = "n%d." % node
    x, y, hmm, env = H.node_inputs(d, node)
    dx, dy = ctx.upload(x), ctx.upload(y)
    pair = ctx.prepare(dx, dy, hmm, env)
    ge = pair.envelope(full=True)
    assert H.same(ge["x_near_start"], d[pfx + "x_near_start"]) and H.same(ge["y_near_end"], d[pfx + "y_near_end"])
    assert H.same(ge["in_envelope"].sum(axis=1).astype(np.int32), d[pfx + "fwd_row_count"])
    assert pair.n_cells == int(d[pfx + "fwd_row_count"].sum())
    info = pair.info
    gt = pair.tables()
    A = x.alph_size
    for k in ("insx", "insy", "rootsubx", "rootsuby"):
        H.assert_log_close(gt[k][..., None], d[pfx + k][..., None], 0.0, (A * x.components + 1) * 4.6e-5, what=cid + ":" + k)
    lp_end = pair.forward()
    lp_start = pair.backward()
    ref_end, ref_start = float(d[pfx + "lp_end"][0]), float(d[pfx + "lp_start"][0])
    assert abs(lp_end - ref_end) <= REL_GOLD * abs(ref_end) + ABS_GOLD, (lp_end, ref_end)
    assert abs(lp_start - ref_start) <= REL_GOLD * abs(ref_start) + ABS_GOLD, (lp_start, ref_start)
    gfwd, gbwd = pair.cells_dense(0), pair.cells_dense(1)
    i, j, v = _sparse(d, pfx, "fwd")
    H.assert_log_close(gfwd[i, j], v, REL_GOLD, ABS_GOLD, what=cid + ":fwd")
    i, j, v = _sparse(d, pfx, "bwd")
    H.assert_log_close(gbwd[i, j], v, REL_GOLD, ABS_GOLD, what=cid + ":bwd")
    # the whole matrix against exact arithmetic
    O.set_exact_lse(True)
    try:
        efwd, elp_end = O.forward(x, y, hmm, env)
        ebwd, elp_start = O.backward(x, y, hmm, env)
    finally:
        O.set_exact_lse(False)
    H.assert_log_close(gfwd, efwd, REL_EXACT, ABS_EXACT, what=cid + ":fwd-exact")
    H.assert_log_close(gbwd, ebwd, REL_EXACT, ABS_EXACT, what=cid + ":bwd-exact")
    assert abs(lp_end - elp_end) <= REL_EXACT * abs(elp_end) + ABS_EXACT
    assert abs(lp_start - elp_start) <= REL_EXACT * abs(elp_start) + ABS_EXACT
    assert abs(lp_start - lp_end) <= 1e-5 * abs(lp_end) + 1e-3
    # these nodes must really exercise what the leaf-level benchmark does not
    in_deg = max(np.bincount(x.trans_dest).max(), np.bincount(y.trans_dest).max())
    assert in_deg > 1 and info.max_diag_width > 32, (in_deg, info.max_diag_width)
    if "lg16" in cid:
        assert info.max_diag_width > 96, info.max_diag_width
