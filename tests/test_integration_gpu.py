"""The compiled reference-side binding (integration/forward_b200.cpp): the reference's UNMODIFIED test mains
(t/testforward.cpp, t/testbackward.cpp, t/testnullforward.cpp) and its `historian` main, built against the reference's own
headers with the DPMatrix / ForwardMatrix / BackwardMatrix constructors bound to libhistorian_b200.so, must reproduce the
reference's own golden outputs (Makefile:247-262, 314-322 of the reference; the files are copied to tests/golden/ref_t)
through the GPU: text identical, numbers within 1e-4 relative (the reference's table log-sum-exp against the device's
exact sums; 6 significant digits are printed).  The binaries are built by `make -C integration` where /root/reference
exists and travel under oracle/_ref/bin."""
import os
import re
import subprocess

import pytest

import helpers as H

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "bin")
DATA = os.path.join(H.GOLDEN_DIR, "ref_t")

CASES = [
    ("b200_testforward", "-all -matrix data/testforward.id100.len2.fa data/testforward.nosub.json 1", "testforward.id100.len2.nosub.out"),
    ("b200_testforward", "-hubs -best data/testforward.len2.fa data/testforward.nosub.json 1", "testforward.len2.nosub.best.out"),
    ("b200_testforward", "-hubs -best data/testforward.len2.fa data/testforward.jukescantor.json 1", "testforward.len2.jc.best.out"),
    ("b200_testforward", "-hubs -best data/testforward.len2-4.fa data/testforward.jukescantor.json .1 .01", "testforward.len2-4.xdel.out"),
    ("b200_testforward", "-hubs -best data/testforward.len2-4.fa data/testforward.jukescantor.json .01 1", "testforward.len2-4.yins.out"),
    ("b200_testforward", "-all 10 data/testforward.len2-4.fa data/testforward.jukescantor.json .1", "testforward.len2-4.n10.all.out"),
    ("b200_testforward", "-hubs 10 data/testforward.len2-4.fa data/testforward.jukescantor.json .1", "testforward.len2-4.n10.hubs.out"),
    ("b200_testnullforward", "data/testforward.nosub.json 1", "testnullforward.nosub.out"),
    ("b200_testbackward", "data/testforward.len2.fa data/testforward.jukescantor.json 1", "testbackward.len2.out"),
    ("b200_testbackward", "data/testforward.len2-4.fa data/testforward.jukescantor.json 1", "testbackward.len2-4.out"),
    ("b200_historian", "recon -careful -norefine -output fasta -model data/testcount.jukescantor.json -guide data/testcount.fa -tree data/testcount.nh",
     "testcount.historian.fa"),
    ("b200_historian", "recon -careful -norefine -output fasta -model data/testnj.jukescantor.json -nexus data/testnexus.nex", "testnexus.hist.fa"),
    ("b200_historian", "recon -careful -norefine -output fasta -profsamples 100 -guide data/PF16593.testspan.fa -model data/testamino.json "
                       "-tree data/PF16593.testspan.testnj.nh -band 10", "PF16593.testspan.testnj.historian.fa"),
]
NUM = re.compile(r"^-?(?:\d+\.?\d*|\.\d+)(?:[eE][-+]?\d+)?$")


def tokens(text):
    return re.findall(r"-?(?:\d+\.?\d*|\.\d+)(?:[eE][-+]?\d+)?|[^\s\d.-]+|[-.]", text)


def same_output(got, want, rel=1e-4, abs_=2e-4):
    g, w = tokens(got), tokens(want)
    if len(g) != len(w):
        return False, "token counts differ: %d vs %d" % (len(g), len(w))
    for k, (a, b) in enumerate(zip(g, w)):
        if a == b:
            continue
        if NUM.match(a) and NUM.match(b):
            x, y = float(a), float(b)
            if abs(x - y) <= abs_ + rel * max(abs(x), abs(y)):
                continue
        return False, "token %d: %r vs %r" % (k, a, b)
    return True, ""


def test_output_comparison_rule():
    assert same_output("lp -12.34561 x", "lp -12.3456 x")[0]
    assert not same_output("lp -12.36 x", "lp -12.3456 x")[0]
    assert not same_output("a b", "a c")[0] and not same_output("a", "a b")[0]


@pytest.mark.parametrize("prog,args,golden", CASES, ids=[c[2] for c in CASES])
def test_reference_main_through_the_binding(tmp_path, prog, args, golden):
    exe = os.path.join(BIN, prog)
    if not os.path.exists(exe):
        pytest.skip("%s not built (make -C integration, needs /root/reference)" % prog)
    (tmp_path / "data").symlink_to(DATA)
    r = subprocess.run([exe] + args.split(), cwd=tmp_path, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    ok, why = same_output(r.stdout.decode(), open(os.path.join(DATA, golden)).read())
    assert ok, why


def test_the_binding_really_calls_the_library(tmp_path):
    """Without a usable device the bound constructors must fail loudly (no CPU fallback inside the binding)."""
    exe = os.path.join(BIN, "b200_testbackward")
    if not os.path.exists(exe):
        pytest.skip("not built")
    (tmp_path / "data").symlink_to(DATA)
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([exe, "data/testforward.len2.fa", "data/testforward.jukescantor.json", "1"], cwd=tmp_path,
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=env, timeout=120)
    assert r.returncode != 0 and b"historian_b200" in r.stderr
