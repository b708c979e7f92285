"""Shared helpers for the parity tests."""
import glob
import os

import numpy as np

from historian_b200 import descs as D

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


# fixtures of large matrices, dumped sparsely (REF_DUMP_SPARSE, oracle/ref_dump.cpp): every stride-th in-envelope cell
SPARSE = ("gp120_nodes.npz", "lg16x300_n14.npz")


def golden_files(sparse=False):
    return sorted(f for f in glob.glob(os.path.join(GOLDEN_DIR, "*.npz"))
                  if not os.path.basename(f).startswith(("sumprod_", "fbcounts_", "quickalign_", "distmat_", "guide_", "elimcounts_")) and (os.path.basename(f) in SPARSE) == sparse)


def load_golden(path):
    z = np.load(path)
    return {k: z[k] for k in z.files}


def golden_cases(sparse=False):
    """[(id, path, node)] for every dumped node of every fixture."""
    cases = []
    for f in golden_files(sparse):
        z = np.load(f)
        for n in z["dumped_nodes"]:
            cases.append(("%s-n%d" % (os.path.basename(f)[:-4], int(n)), f, int(n)))
    return cases


def node_inputs(d, node):
    pfx = "n%d." % node
    return (D.profile_from_dump(d, pfx + "x."), D.profile_from_dump(d, pfx + "y."),
            D.hmm_from_dump(d, pfx + "hmm."), D.env_from_dump(d, pfx + "env."))


def same(a, b):
    a, b = np.asarray(a), np.asarray(b)
    return a.shape == b.shape and bool(np.all((a == b) | ((a != a) & (b != b))))


def assert_log_close(got, ref, rel, abs_, floor=80.0, what=""):
    """Compare natural-log arrays whose last axis holds the 5 states of a cell.

    |got-ref| <= abs_ + rel*|ref| wherever both are finite.  ref == -inf requires got == -inf.
    A state lying more than `floor` nats below the largest state of the SAME cell is exempt
    from the precision check and may come back as -inf: the five FP32 values of a cell share
    one exponent, so the dynamic range INSIDE a cell is 2^-126 ~ e^-87 (denormal below that);
    such a state carries < e^-80 of the cell's probability and cannot influence any result."""
    got, ref = np.asarray(got, float), np.asarray(ref, float)
    assert got.shape == ref.shape, (what, got.shape, ref.shape)
    ref_inf, got_inf = np.isneginf(ref), np.isneginf(got)
    assert not np.isnan(got).any(), what + ": NaN in result"
    bad = ref_inf & ~got_inf
    assert not bad.any(), "%s: %d entries finite where the reference has -inf (first %s)" % (what, bad.sum(), np.argwhere(bad)[:3].tolist())
    cellmax = np.max(np.where(np.isfinite(ref), ref, -np.inf), axis=-1, keepdims=True)
    with np.errstate(invalid="ignore"):
        deep = np.broadcast_to(ref < cellmax - floor, ref.shape)
    lost = got_inf & ~ref_inf
    assert not (lost & ~deep).any(), "%s: %d entries lost (-inf) that are within %g nats of their cell max" % (what, (lost & ~deep).sum(), floor)
    with np.errstate(invalid="ignore"):
        assert not (deep & ~got_inf & (got > cellmax - floor + 1)).any(), what + ": a deep entry came back large"
    fin = ~ref_inf & ~got_inf & ~deep
    if fin.any():
        err = np.abs(got[fin] - ref[fin])
        tol = abs_ + rel * np.abs(ref[fin])
        worst = np.argmax(err - tol)
        assert (err <= tol).all(), "%s: max |diff| %.3g (tol %.3g) at value %.6f vs %.6f; %d/%d out of tolerance" % (
            what, err[worst], tol[worst], got[fin][worst], ref[fin][worst], (err > tol).sum(), err.size)
    return float(np.max(np.abs(got[fin] - ref[fin]))) if fin.any() else 0.0
