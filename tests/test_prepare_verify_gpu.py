"""The integer side of the DPMatrix constructors is built on the device (profile structure images, band keys, monotone hull,
anti-diagonal layout: csrc/hist_core.cu pair_layout_kernel).  With HIST_B200_VERIFY_PREP=1 the library reads every pair's
keys, edge flags, state flags and layout records back and compares them with a plain host computation of the same integers
(forward.cpp:36-42, 58-65).  Run here on leaf x leaf merges and on every internal (DAG x DAG) merge of a whole-tree
reconstruction, deterministic and sampled; a child process, because the switch is read once per process."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(args):
    env = dict(os.environ, HIST_B200_VERIFY_PREP="1")
    r = subprocess.run([sys.executable] + args, cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("[verify_prep]")]
    assert not any("MISMATCH" in l or "device" in l for l in lines), "\n".join(l for l in lines if "ok" not in l)[:3000]
    return lines


def test_leaf_merges():
    lines = run([os.path.join("tools", "dbg_prep.py"), "8"])
    assert sum(l.endswith(": ok") for l in lines) == 3 * 8


@pytest.mark.parametrize("extra", [[], ["--samples"]], ids=["best_trace_profiles", "sampled_profiles"])
def test_every_merge_of_a_reconstruction(extra):
    lines = run([os.path.join("tools", "recon_native_bench.py"), "16", "300"] + extra)
    assert sum(l.endswith(": ok") for l in lines) >= 2 * 15   # two repetitions of 15 merges (band retries add more)
