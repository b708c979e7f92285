"""Codon alphabets (BASELINE config 5, ECMrest: 61 sense codons): Profile::leftMultiply and the DPMatrix-constructor
tables on the tensor cores (historian_b200/csrc/hist_codon.cu, tcgen05.mma kind::tf32 with a 3xTF32 split, FP32
accumulators in tensor memory) against the exact-arithmetic oracle, against the reference's dumps (tests/golden/
ecmrest8.npz, through the generic fixture tests) and against the CUDA-core kernel it replaces (same library, run in a
child process with HIST_B200_NO_TENSOR=1)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import helpers as H
import oracle_lib as O
import synth
from historian_b200 import descs as D, model as M

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _pair_inputs(seed=5, lx=130, ly=150):
    rng = np.random.default_rng(seed)
    rm = M.RateModel.preset("ECMrest")
    x, y, (r1, r2) = synth.leaf_pair(rng, rm.alphabet, lx, ly, identity=0.6)
    return x, y, M.pair_hmm(rm, 0.2, 0.35), D.Envelope.from_guide_rows(r1, r2, 10)


def test_tensor_tables_match_exact_oracle(ctx):
    """Leaf x leaf (one-hot rows) and DAG x DAG (the internal nodes of the ECMrest fixture: dense rows): every table of the
    constructor within 5e-6 of exact arithmetic, Forward/Backward within the usual fill tolerance."""
    cases = [_pair_inputs()]
    d = H.load_golden(os.path.join(H.GOLDEN_DIR, "ecmrest8.npz"))
    for node in (6, 13, 14):
        cases.append(H.node_inputs(d, node))
    for x, y, hmm, env in cases:
        assert x.alph_size == 61 and x.components == 1
        dx, dy = ctx.upload(x), ctx.upload(y)
        pair = ctx.prepare(dx, dy, hmm, env)
        O.set_exact_lse(True)
        try:
            et = O.pair_tables(x, y, hmm, env)
            ef, elp = O.forward(x, y, hmm, env)
        finally:
            O.set_exact_lse(False)
        gt = pair.tables()
        worst = 0.0
        for k in ("subx", "suby", "insx", "insy", "rootsubx", "rootsuby"):
            worst = max(worst, H.assert_log_close(gt[k][..., None], et[k][..., None], 0.0, 5e-6, what=k))
        lp = pair.forward()
        assert abs(lp - elp) <= 1e-5 * abs(elp) + 2e-4, (lp, elp)
        H.assert_log_close(pair.cells_dense(0), ef, 1e-5, 2e-4, what="fwd")
        assert worst > 0.0   # not bit-identical to FP64: the products really went through FP32 tensor arithmetic


CHILD = r"""
import json, sys
sys.path.insert(0, %r); sys.path.insert(0, %r)
import numpy as np
from test_codon_gpu import _pair_inputs
from historian_b200 import capi
ctx = capi.Context(0)
x, y, hmm, env = _pair_inputs()
pair = ctx.prepare(ctx.upload(x), ctx.upload(y), hmm, env)
t = pair.tables()
print(json.dumps({k: np.where(np.isfinite(t[k]), t[k], -1e300).ravel().tolist() for k in ("subx", "rootsuby", "insx")}))
"""


def test_tensor_path_differs_from_and_agrees_with_cuda_core_path(ctx):
    x, y, hmm, env = _pair_inputs()
    pair = ctx.prepare(ctx.upload(x), ctx.upload(y), hmm, env)
    t = pair.tables()
    env2 = dict(os.environ, HIST_B200_NO_TENSOR="1")
    out = subprocess.run([sys.executable, "-c", CHILD % (ROOT, os.path.join(ROOT, "tests"))], stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                         env=env2, check=True, timeout=600).stdout.decode()
    ref = json.loads(out.strip().splitlines()[-1])
    differs = False
    for k, v in ref.items():
        a = np.where(np.isfinite(t[k]), t[k], -1e300).ravel()
        b = np.array(v)
        assert a.shape == b.shape
        assert np.array_equal(a < -1e299, b < -1e299)
        fin = a > -1e299
        assert np.max(np.abs(a[fin] - b[fin])) <= 5e-6, (k, np.max(np.abs(a[fin] - b[fin])))
        differs = differs or bool(np.any(a[fin] != b[fin]))
    assert differs, "the two constructors gave bit-identical tables: the tensor path did not run"
