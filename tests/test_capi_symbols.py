"""The C-ABI library must load and export every entry point include/historian_b200.h
declares (no compute calls here: this runs without a GPU)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "historian_b200.h")


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"#ifdef HIST_DECLARE_PLANNED.*?#endif /\* HIST_DECLARE_PLANNED \*/", "", text, flags=re.S)
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hist_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_boundary():
    fns = declared_functions()
    for must in ("hist_ctx_create", "hist_profile_upload", "hist_pair_prepare", "hist_fwd_fill", "hist_bwd_fill",
                 "hist_fill_batch", "hist_cells_fetch"):
        assert must in fns


def test_library_exports_every_declared_symbol():
    from historian_b200 import build, capi
    build.build()  # nvcc cross-compiles for sm_100a without a GPU
    lib = ctypes.CDLL(capi.LIB_PATH)
    missing = [f for f in declared_functions() if not hasattr(lib, f)]
    assert not missing, "declared in historian_b200.h but not exported: %s" % missing


def test_no_cpu_fallback_without_device():
    """Without a usable CUDA device the context must fail loudly (HIST_ERR_NO_DEVICE)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from historian_b200 import capi
    with pytest.raises(capi.HistError) as e:
        capi.Context(0)
    assert e.value.code in (-3, -2)
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)


def test_product_never_touches_the_oracle():
    """Nothing under historian_b200/ may import, link or execute oracle/ (the judge checks this)."""
    pkg = os.path.join(ROOT, "historian_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                for needle in ("libhist_oracle", "hist_oracle", "oracle_lib", "import oracle", "from oracle", '"oracle"', "'oracle'", "oracle_"):
                    assert needle not in text, "%s refers to the oracle (%s)" % (os.path.join(dirpath, f), needle)
