"""`historian fit` on a reconstructed alignment (historian_b200/em.py: EventCounts.optimize = EventCounts::optimize,
model.cpp:1022-1059; fit = Reconstructor::fit, recon.cpp:1385-1408) against the reference BINARY run at test time on the
same input: per-iteration log-likelihood and log-prior, and the fitted model.

  CPU  : the E-step of the loop is the CPU oracle with the reference's table log-sum-exp -> 6 significant digits.
  GPU  : the E-step is the device path (exact log-sum-exp in the column likelihoods of a 2-component mixture) -> 1e-4.
"""
import json
import os
import re
import subprocess
import tempfile

import numpy as np
import pytest

import helpers as H
import oracle_lib as O
from historian_b200 import descs as D, em, model as M, simulator as S

REF_BIN = os.path.join(os.path.dirname(H.GOLDEN_DIR), "..", "oracle", "_ref", "bin", "historian")


def _family(n_leaves=16, root_len=120, seed=3):
    """A full history (ancestral rows included) drawn from a model that is NOT the starting point of the fit: unequal
    base frequencies, transitions four times as frequent as transversions, more and longer indels."""
    tree = S.Tree.balanced(n_leaves, 0.3)
    freq = {"a": 0.35, "c": 0.15, "g": 0.2, "t": 0.3}
    ts = {("a", "g"), ("g", "a"), ("c", "t"), ("t", "c")}
    truth = M.RateModel.read({"alphabet": "acgt", "insrate": 0.03, "delrate": 0.04, "insextprob": 0.5, "delextprob": 0.4, "rootprob": freq,
                              "subrate": {i: {j: (2.0 if (i, j) in ts else 0.5) * freq[j] for j in "acgt" if j != i} for i in "acgt"}})
    gapped, _ = S.simulate_family(truth, tree, root_len, seed)
    return tree, gapped


def _newick_with_internal_names(t):
    def rec(v):
        s = ("(" + ",".join(rec(c) for c in t.children[v]) + ")" if t.children[v] else "") + t.names[v]
        return s + (":%g" % t.branch_length[v] if t.parent[v] >= 0 else "")
    return rec(t.root) + ";\n"


def _reference_fit(tree, gapped, extra):
    with tempfile.TemporaryDirectory() as tmp:
        open(os.path.join(tmp, "tree.nh"), "w").write(_newick_with_internal_names(tree))
        with open(os.path.join(tmp, "recon.fa"), "w") as f:
            for v in range(tree.n_nodes):
                f.write(">%s\n%s\n" % (tree.names[v], bytes(gapped[v]).decode()))
        r = subprocess.run([REF_BIN, "fit", "-recon", "recon.fa", "-tree", "tree.nh", "-preset", "jc", "-gamma", "2", "-v1"] + extra,
                           stdout=subprocess.PIPE, stderr=subprocess.PIPE, cwd=tmp, check=True)
    its = []
    for line in r.stderr.decode().splitlines():
        m = re.search(r"EM iteration #(\d+): log-likelihood \((-?[\d.eE+-]+)\) \+ log-prior \((-?[\d.eE+-]+)\)", line)
        if m:
            its.append((float(m.group(2)), float(m.group(3))))
            continue
        m = re.search(r"EM iteration #(\d+): log-likelihood = (-?[\d.eE+-]+)", line)
        if m:
            its.append((float(m.group(2)), 0.0))
    return its, json.loads(r.stdout.decode())


def _check(model, lps, ref_its, ref_model, rel):
    assert len(lps) == len(ref_its) and len(lps) >= 3, (lps, ref_its)
    for (lp, pr), (rlp, rpr) in zip(lps, ref_its):
        assert abs(lp - rlp) <= rel * abs(rlp) + 1e-6, (lp, rlp)
        assert abs(pr - rpr) <= max(rel, 1e-5) * abs(rpr) + 1e-5, (pr, rpr)
    fitted = M.RateModel.read(ref_model)
    tol = max(10 * rel, 2e-5)
    for a, b in ((model.insRate, fitted.insRate), (model.delRate, fitted.delRate), (model.insExtProb, fitted.insExtProb),
                 (model.delExtProb, fitted.delExtProb)):
        assert abs(a - b) <= tol * abs(b), (a, b)
    assert np.allclose(model.cptWeight, fitted.cptWeight, rtol=tol)
    for c in range(model.components):
        assert np.allclose(np.asarray(model.insProb[c]) / np.sum(model.insProb[c]), fitted.insProb[c], rtol=tol, atol=1e-9)
        assert np.allclose(model.subRate[c], fitted.subRate[c], rtol=tol, atol=1e-9)


def _oracle_counter(tree, cols):
    topo = D.TreeTopology(tree.parent, tree.branch_length)

    def counter(model):
        spm = D.SumProdModel.from_rate_model(model)
        counts = O.sumprod_columns(spm, topo, cols, flags=O.SP_COUNTS)["counts"]
        counts[:7] += em.indel_counts(cols, topo.parent, topo.branch_length, model.insRate, model.delRate, model.insExtProb, model.delExtProb)
        root, sub = O.sumprod_transform_counts(spm, counts)
        return em.EventCounts(counts[:7], root, sub)
    return counter


@pytest.mark.parametrize("laplace", [True, False])
def test_fit_loop_matches_reference_binary_cpu(laplace):
    if not os.path.exists(REF_BIN):
        pytest.skip("oracle/_ref/bin/historian not built")
    tree, gapped = _family()
    # -mininc -1: keep iterating whatever the likelihood does (the reference stops as soon as it fails to improve by 0.1 %)
    ref_its, ref_model = _reference_fit(tree, gapped, ["-maxiter", "6", "-mininc", "-1"] + ([] if laplace else ["-nolaplace"]))
    cols = np.ascontiguousarray(gapped.T)
    model = M.RateModel.preset("jc").gamma(2, 1.0)
    model, lps = em.fit(None, model, None, cols, max_iter=6, min_improvement=-1.0, prior="laplace" if laplace else None,
                        counter=_oracle_counter(tree, cols))
    _check(model, lps, ref_its, ref_model, 2e-6)


def test_m_step_is_the_expected_log_likelihood_maximiser():
    """optimize() must not lower EventCounts::expectedLogLikelihood (model.cpp:1083-1104), and small moves away lower it."""
    tree, gapped = _family(8, 60, 5)
    cols = np.ascontiguousarray(gapped.T)
    model = M.RateModel.preset("jc").gamma(2, 1.0)
    ev = _oracle_counter(tree, cols)(model)
    before = ev.expected_log_likelihood(model)
    ev.optimize(model)
    best = ev.expected_log_likelihood(model)
    assert best >= before
    rng = np.random.default_rng(0)
    for _ in range(5):
        m2 = M.RateModel.preset("jc").gamma(2, 1.0)
        ev.optimize(m2)
        m2.insRate *= 1 + 0.05 * rng.standard_normal()
        m2.subRate[0] = m2.subRate[0] * (1 + 0.05 * rng.standard_normal((4, 4)))
        m2.subRate[0][np.diag_indices(4)] = 0
        m2.subRate[0][np.diag_indices(4)] = -m2.subRate[0].sum(axis=1)
        assert ev.expected_log_likelihood(m2) <= best + 1e-9


@pytest.mark.gpu
def test_fit_loop_matches_reference_binary_device(ctx):
    if not os.path.exists(REF_BIN):
        pytest.skip("oracle/_ref/bin/historian not built")
    tree, gapped = _family()
    ref_its, ref_model = _reference_fit(tree, gapped, ["-maxiter", "6", "-mininc", "-1"])
    cols = np.ascontiguousarray(gapped.T)
    model = M.RateModel.preset("jc").gamma(2, 1.0)
    model, lps = em.fit(ctx, model, D.TreeTopology(tree.parent, tree.branch_length), cols, max_iter=6, min_improvement=-1.0)
    _check(model, lps, ref_its, ref_model, 1e-4)
