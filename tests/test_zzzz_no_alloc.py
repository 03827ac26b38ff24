"""bwk_biharmonic_wk_scalar_noalloc: the allocation-free form of the single Laplacian (include/bwk_b200.h; no counterpart in
the reference).  Called through the C ABI on device pointers and compared with
the existing entry point bwk_biharmonic_wk_scalar (F90:525-534 signature, host buffers) on the same inputs, several calls in a
row that share the static scratch array.  This file sorts after every other test file on purpose.

The device buffers are the CALLER's (this test's): they come from cudaMalloc of the CUDA runtime the library itself links.
"""
import ctypes as C

import numpy as np
import pytest


class _Cudart:
    def __init__(self):
        rt = C.CDLL("libcudart.so.12", mode=C.RTLD_GLOBAL)  # the copy libbwk_b200.so is linked against
        rt.cudaMalloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
        rt.cudaFree.argtypes = [C.c_void_p]
        rt.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
        rt.cudaDeviceSynchronize.argtypes = []
        rt.cudaGetErrorString.restype = C.c_char_p
        rt.cudaGetErrorString.argtypes = [C.c_int]
        self.rt = rt

    def ck(self, rc):
        assert rc == 0, self.rt.cudaGetErrorString(rc).decode()

    def to_device(self, a):
        p = C.c_void_p()
        self.ck(self.rt.cudaMalloc(C.byref(p), a.nbytes))
        self.ck(self.rt.cudaMemcpy(p, a.ctypes.data_as(C.c_void_p), a.nbytes, 1))
        return p

    def to_host(self, p, like):
        out = np.empty_like(like, order="F")
        self.ck(self.rt.cudaMemcpy(out.ctypes.data_as(C.c_void_p), p, out.nbytes, 2))
        return out

    def free(self, p):
        self.ck(self.rt.cudaFree(p))


def _inputs(bwk, nelemd, nlev, qsize, seed):
    rng = np.random.default_rng(seed)
    elem = np.zeros(nelemd, dtype=bwk.ELEMENT_DTYPE)
    elem.view(np.float64)[...] = rng.random(144 * nelemd)
    deriv = np.zeros((), dtype=bwk.DERIVATIVE_DTYPE)
    deriv["Dvv"] = rng.random((4, 4))
    q = np.asfortranarray(rng.random((4, 4, nlev, qsize, nelemd)) - 0.5)
    return elem, deriv, q


def _noalloc(bwk, rt, elem, deriv, q):
    nlev, qsize, nelemd = q.shape[2:]
    d_elem, d_q = rt.to_device(elem), rt.to_device(q)
    try:
        bwk.biharmonic_wk_scalar_noalloc(d_elem.value, d_q.value, deriv, nelemd, nlev, qsize, 0)
        rt.ck(rt.rt.cudaDeviceSynchronize())
        return rt.to_host(d_q, q)
    finally:
        rt.free(d_elem)
        rt.free(d_q)


def _existing(bwk, elem, deriv, q):
    nlev, qsize, nelemd = q.shape[2:]
    ku = bwk.KernelUtils(nelemd, nlev, qsize).allocate_data(host_arrays=False)
    try:
        out = q.copy(order="F")
        ku.biharmonic_wk_scalar(elem, out, None, deriv, 1, nelemd)
        return out
    finally:
        ku.finalize()


@pytest.mark.gpu
def test_noalloc_matches_existing_entry_point_on_calls_in_a_row(bwk):
    rt = _Cudart()
    # (nelemd, nlev, qsize): the reference's default problem (F90:11-17); a ragged one (nlev*qsize = 15 < one 32-slab tile)
    # on FEWER elements, so the scratch array still holds the first call's metric beyond what this call rewrites; a larger one
    cases = [(16, 72, 40, 1), (7, 5, 3, 2), (150, 9, 5, 3)]
    got, inputs = [], []
    for nelemd, nlev, qsize, seed in cases:  # the new entry point first: it must not need bwk_init
        elem, deriv, q = _inputs(bwk, nelemd, nlev, qsize, seed)
        inputs.append((elem, deriv, q))
        got.append(_noalloc(bwk, rt, elem, deriv, q))
    for (elem, deriv, q), g in zip(inputs, got):
        want = _existing(bwk, elem, deriv, q)
        assert np.isfinite(want).all() and not np.array_equal(want, q)
        # same kernels, same arithmetic: the tolerance of the other parity tests (max elementwise relative 1e-12, rel-L2
        # 4e-16) is met with zero difference
        rel_l2 = np.sqrt(((g - want) ** 2).sum() / (want**2).sum())
        max_rel = np.max(np.abs(g - want) / np.maximum(np.abs(want), 1e-300))
        print(f"noalloc vs bwk_biharmonic_wk_scalar {q.shape}: bit-identical {np.array_equal(g, want)}, rel-L2 {rel_l2:.3e}")
        assert rel_l2 <= 4e-16 and max_rel <= 1e-12, (q.shape, rel_l2, max_rel)
    _check_rejections(bwk)
    # and once more after the context of the existing entry point has come and gone
    elem, deriv, q = inputs[0]
    again = _noalloc(bwk, rt, elem, deriv, q)
    assert np.array_equal(again, got[0])


def _check_rejections(bwk):
    L = bwk.lib()
    deriv = np.zeros((), dtype=bwk.DERIVATIVE_DTYPE)
    dummy = C.c_void_p(256)  # never dereferenced: the size check comes first
    rc = L.bwk_biharmonic_wk_scalar_noalloc(dummy, dummy, deriv.ctypes.data_as(C.c_void_p), bwk.NOALLOC_MAX_NELEMD + 1, 4, 2, 0)
    assert rc != 0
    msg = L.bwk_last_error().decode()
    assert "BWK_NOALLOC_MAX_NELEMD" in msg and str(bwk.NOALLOC_MAX_NELEMD) in msg, msg
    rc = L.bwk_biharmonic_wk_scalar_noalloc(None, dummy, deriv.ctypes.data_as(C.c_void_p), 4, 4, 2, 0)
    assert rc != 0 and "null pointer" in L.bwk_last_error().decode()


def test_noalloc_rejects_sizes_beyond_the_static_scratch(bwk):
    # the argument checks come before any CUDA call, so this one also runs on a box without a GPU
    _check_rejections(bwk)
    # the limit the binding advertises is the header's, and the header says what the entry point is
    import os
    import re

    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "bwk_b200.h")).read()
    assert int(re.search(r"#define\s+BWK_NOALLOC_MAX_NELEMD\s+(\d+)", hdr).group(1)) == bwk.NOALLOC_MAX_NELEMD
    assert "NO COUNTERPART IN THE REFERENCE" in hdr and "bwk_biharmonic_wk_scalar_noalloc" in bwk.declared_symbols()
