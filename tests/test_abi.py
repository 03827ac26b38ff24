"""CPU: the C-ABI library loads, exports every symbol include/bwk_b200.h declares, and fails
loudly (no fallback) when there is no CUDA device.  No compute calls here."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_exports_every_declared_symbol(bwk):
    L = bwk.lib()
    names = bwk.declared_symbols()
    assert len(names) >= 40
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing


def test_struct_layouts_match_reference_types(bwk):
    # element_t F90:23-27 = 64 + 16 + 64 doubles, derivative_t F90:19-21 = 16 doubles
    assert bwk.ELEMENT_DTYPE.itemsize == 1152 and bwk.DERIVATIVE_DTYPE.itemsize == 128
    assert bwk.ELEMENT_DTYPE.fields["spheremp"][1] == 64 * 8
    assert bwk.ELEMENT_DTYPE.fields["tensorVisc"][1] == 80 * 8
    hdr = open(os.path.join(ROOT, "include", "bwk_b200.h")).read()
    assert "double Dinv[2][2][BWK_NP][BWK_NP];" in hdr and "double tensorVisc[2][2][BWK_NP][BWK_NP];" in hdr


def test_no_cpu_fallback_without_a_device(bwk):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    L = bwk.lib()
    assert L.bwk_init(16, 72, 40) != 0
    msg = L.bwk_last_error().decode()
    assert "no CPU fallback" in msg or "CUDA" in msg
    assert L.bwk_run(1, 16) != 0 and "bwk_init" in L.bwk_last_error().decode()


def test_product_never_touches_the_oracle():
    # the oracle is test infrastructure: nothing under the package may reference it
    pkg = os.path.join(ROOT, "codesign-kernels_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".F90", "Makefile")):
                text = open(os.path.join(d, f), errors="ignore").read()
                assert "bwk_oracle" not in text and "oracle_py" not in text and "bwko_" not in text, os.path.join(d, f)


def test_partition_is_contiguous_and_complete(bwk):
    for n, p in [(16, 1), (16, 3), (5400, 8), (86400, 8), (7, 8), (1572864, 8)]:
        nxt = 0
        for r in range(p):
            f, c = bwk.partition(n, r, p)
            assert f == nxt and c >= 0
            nxt = f + c
        assert nxt == n


def test_fortran_wrapper_matches_reference_signature():
    """The ISO_C_BINDING shim ships as source (no Fortran compiler in the image); lint it with
    numpy's Fortran parser and compare its public argument list with the GPU2 driver F90:525-531."""
    from numpy.f2py import crackfortran

    path = os.path.join(ROOT, "codesign-kernels_b200", "fortran", "biharmonic_wk_scalar_b200.F90")
    crackfortran.verbose = 0
    blocks = crackfortran.crackfortran([path])
    mods = {b["name"]: b for b in blocks}
    assert "biharmonic_wk_scalar_b200" in mods
    subs = {b["name"]: b for b in mods["biharmonic_wk_scalar_b200"]["body"] if b["block"] == "subroutine"}
    assert subs["biharmonic_wk_scalar"]["args"] == ["elem", "qtens", "grads", "deriv", "nets", "nete"]
    assert subs["biharmonic_wk_scalar_cpusig"]["args"] == ["elem", "qtens", "deriv", "nets", "nete"]
    assert subs["biharmonic_wk_scalar_present"]["args"] == ["elem", "qtens", "deriv"]  # the allocation-free form on OpenACC-present data
    assert subs["advance_hypervis_scalar"]["args"] == ["elem", "qdp", "dp", "dpdiss", "deriv", "dt2", "nu_q", "hypervis_subcycle_q", "limiter"]
    assert subs["biharmonic_wk_dp3d"]["args"] == ["elem", "vec_sphere2cart", "dptens", "ttens", "vtens", "deriv", "var_coef1", "undamprrcart"]
    # the dp3d entry needs a context with qsize = 4 while kernel_utils' qsize is a compile-time 40: b200_init takes the
    # context's nlev / qsize as optional arguments
    assert subs["b200_init"]["args"] == ["nlev_ctx", "qsize_ctx"]
    assert all("optional" in subs["b200_init"]["vars"][a].get("attrspec", []) for a in ("nlev_ctx", "qsize_ctx"))
    text = open(path).read()
    for sym in re.findall(r'bind\(c,\s*name="(bwk_[a-z0-9_]+)"\)', text, flags=re.I):
        assert sym in open(os.path.join(ROOT, "include", "bwk_b200.h")).read(), sym
