"""Refinement stage: Refiner (refiner.cpp) = conditional position weight matrices
(hist_sumprod_excluded_posteriors), branch Viterbi alignment (hist_branch_align), merge, log-likelihood test.

 * the excluded posteriors against the E / G messages the library already exposes for parity (hist_sumprod_debug_column,
   themselves pinned on the reference's SumProduct, tests/test_sumprod_gpu.py), combined as sumprod.cpp:219-250 does;
 * the branch alignment against oracle/branch_align.py (the reference's loops in plain Python, exact log-sum-exp);
 * the whole refinement against the reference BINARY: `historian recon ... -refine` on BASELINE configs 1 and 2, started
   from the binary's own unrefined reconstruction, must end in the identical alignment."""
import os
import subprocess
import sys

import numpy as np
import pytest

import helpers as H
import oracle_lib as O
from historian_b200 import descs as D, model as M, refine as RF, simulator as S
from test_recon_gpu import _read_fasta, REF_BIN
from test_configs_gpu import CONFIGS

sys.path.insert(0, O.ORACLE_DIR)
from branch_align import branch_align as oracle_branch_align   # noqa: E402

pytestmark = pytest.mark.gpu


def _history(name, extra):
    gf, tf = CONFIGS[name]
    out = subprocess.run([REF_BIN, "recon", "-guide", gf, "-tree", tf, "-output", "fasta", "-profmaxstates", "1"] + extra,
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True).stdout.decode()
    tree = S.Tree.from_newick(open(tf).read())
    rows = _read_fasta(out)
    assert len(rows) == tree.n_nodes
    return tree, np.array([np.frombuffer(r.encode(), np.uint8) for _, r in rows])


def test_excluded_posteriors_equal_the_messages(ctx):
    tree, rows = _history("pf16593", ["-norefine"])
    lg = M.RateModel.preset("lg")
    sp = ctx.sumprod(D.SumProdModel.from_rate_model(lg), D.TreeTopology(tree.parent, tree.branch_length))
    cols = np.ascontiguousarray(rows.T)
    kids = {v: [c for c in range(tree.n_nodes) if tree.parent[c] == v] for v in range(tree.n_nodes)}
    internal = [v for v in range(tree.n_nodes) if kids[v] and tree.parent[v] >= 0]
    node, leaf = internal[len(internal) // 2], 3
    queries = [(node, kids[node][0]), (node, int(tree.parent[node])), (kids[node][1], node), (leaf, int(tree.parent[leaf])), (tree.n_nodes - 1, -1)]
    got = sp.excluded_posteriors(cols, queries, normalize=False)
    gotn = sp.excluded_posteriors(cols, queries)
    lcw = np.log(np.asarray(lg.cptWeight, float))
    alph = lg.alphabet
    with np.errstate(divide="ignore"):
        for ci in range(0, len(cols), 3):
            m = sp.debug_column(cols[ci])
            for qi, (v, ex) in enumerate(queries):
                ch = chr(cols[ci][v])
                if ch in "-.":
                    assert np.all(np.isneginf(got[qi, ci]))
                    continue
                want = np.zeros((1, len(alph))) + lcw[:, None]
                if ch.lower() in alph:
                    mask = np.full(len(alph), -np.inf)
                    mask[alph.index(ch.lower())] = 0
                    want = want + mask
                for c in kids[v]:
                    if c != ex and chr(cols[ci][c]) not in "-.":
                        want = want + np.log(m["E"][:, c]) + m["logE"][:, c][:, None]
                if tree.parent[v] != ex or ex < 0:
                    want = want + np.log(m["G"][:, v]) + m["logG"][:, v][:, None]
                fin = np.isfinite(want)
                assert np.array_equal(fin, np.isfinite(got[qi, ci])), (ci, qi)
                assert np.allclose(got[qi, ci][fin], want[fin], rtol=0, atol=1e-9), (ci, qi)
                norm = np.log(np.sum(np.exp(want[fin] - want[fin].max()))) + want[fin].max()
                assert np.allclose(gotn[qi, ci][fin], want[fin] - norm, rtol=0, atol=1e-9), (ci, qi)
    sp.close()


def test_branch_alignment_equals_the_oracle(ctx):
    rng = np.random.default_rng(3)
    for trial, (Cc, A, lx, ly, maxd) in enumerate([(1, 4, 7, 9, -1), (2, 4, 12, 10, 2), (1, 20, 15, 15, 3), (1, 4, 1, 6, 0), (1, 4, 5, 0, -1), (1, 4, 0, 0, -1)]):
        sub = rng.dirichlet(np.ones(A), size=(Cc, A))
        with np.errstate(divide="ignore"):
            log_sub, log_ins, lcw = np.log(sub), np.log(rng.dirichlet(np.ones(A), size=Cc)), np.log(rng.dirichlet(np.ones(Cc)))
            lpt = np.log(np.array([[0.8, 0.1, 0.1, 0.9], [0.5, 0.4, 0.1, 0.6], [0.7, 0.0, 0.3, 0.7]]))

            def pwm(n):
                p = np.log(rng.random((n, Cc, A)))
                for k in range(n):   # observed residues: one finite entry per component
                    if rng.random() < 0.5:
                        keep = rng.integers(0, A)
                        p[k, :, np.arange(A) != keep] = -np.inf
                return p
            xp, yp = pwm(lx), pwm(ly)
        # keys of a random old alignment
        kx = np.concatenate([[0], np.cumsum(rng.random(lx) < 0.8)]).astype(np.int32)
        ky = np.concatenate([[0], np.minimum(np.cumsum(rng.random(ly) < 0.8), kx[-1])]).astype(np.int32)
        want_lp, want_path = oracle_branch_align(log_sub.tolist(), log_ins.tolist(), lcw.tolist(), lpt.tolist(), xp.tolist(), yp.tolist(), maxd, kx, ky)
        lp, path = ctx.branch_align(log_sub, log_ins, lcw, lpt, xp, yp, maxd, kx, ky)
        assert abs(lp - want_lp) <= 1e-9 * max(1.0, abs(want_lp)), trial
        assert path.tolist() == want_path, trial
        assert int(np.sum(path & 1)) == lx and int(np.sum(path >> 1)) == ly


@pytest.mark.parametrize("name", ["pf16593", "gp120"])
def test_refinement_equals_the_reference_binary(ctx, name):
    if not os.path.exists(REF_BIN):
        pytest.skip("oracle/_ref/bin/historian not built")
    tree, rows = _history(name, ["-norefine"])
    _, want = _history(name, ["-refine"])
    lg = M.RateModel.preset("lg")
    r = RF.Refiner(ctx, lg, tree)
    try:
        lp0 = r.log_likelihood(rows)
        got, lp = r.refine(rows)
    finally:
        r.close()
    assert lp >= lp0
    assert got.shape == want.shape and np.array_equal(got, want), "%d moves, %d accepted, lp %.6f -> %.6f" % (r.moves, r.accepted, lp0, lp)
    if name == "pf16593":
        assert r.accepted >= 1   # the reference improves this history once (-2073.02 -> -2066.79)
