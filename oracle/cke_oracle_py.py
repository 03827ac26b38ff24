"""TEST INFRASTRUCTURE ONLY -- ctypes view of oracle/cke_oracle.c (the nested_loops loop forms).

Only tests/, __graft_entry__.smoke() and bench.py's / profiles' cpu_baseline legs may import this
module; the product package never does.  Pinned on the reference's own C++ compiled here (oracle/ref_yakl_py.py).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


class _Problem(C.Structure):
    _fields_ = (
        [(n, C.c_int) for n in ("nEdges", "nCells", "nVertLevels", "nvldim", "nAdv")]
        + [(n, C.c_void_p) for n in ("nAdvCellsForEdge", "minLevelCell", "maxLevelCell", "advCellsForEdge")]
        + [(n, C.c_void_p) for n in ("tracerCur", "cellMask", "normalThicknessFlux", "advMaskHighOrder")]
        + [(n, C.c_void_p) for n in ("advCoefs", "advCoefs3rd")]
        + [("coef3rdOrder", C.c_double)]
    )


_libs = {}


def _load(name):
    if name in _libs:
        return _libs[name]
    path = os.path.join(_HERE, name)
    if not os.path.exists(path):
        subprocess.run(["make", "-C", _HERE], check=True, stdout=subprocess.DEVNULL)
    L = C.CDLL(path)
    pp, p = C.POINTER(_Problem), C.c_void_p
    L.ckeo_original_cpu.argtypes = [pp, p]
    L.ckeo_gpu_optimized.argtypes = [pp, p]
    L.ckeo_impl1.argtypes = [pp, p, p]
    L.ckeo_impl2.argtypes = [pp, p, p]
    L.ckeo_check.argtypes = [p, p, C.c_int, C.c_int, C.c_int, C.c_double, C.POINTER(C.c_double)]
    L.ckeo_check.restype = C.c_int64
    L.ckeo_original_cpu_omp.argtypes = [pp, p, C.c_int]
    _libs[name] = L
    return L


def _problem(prob, tracer=None):
    """prob: any object with nested.py's NestedProblem attributes.  Returns (struct, keepalive)."""
    f = lambda x, dt: np.asfortranarray(x, dtype=dt)  # noqa: E731
    keep = dict(
        nAdvCellsForEdge=f(prob.nAdvCellsForEdge, np.int32), minLevelCell=f(prob.minLevelCell, np.int32),
        maxLevelCell=f(prob.maxLevelCell, np.int32), advCellsForEdge=f(prob.advCellsForEdge, np.int32),
        tracerCur=f(prob.tracerCur if tracer is None else tracer, np.float64), cellMask=f(prob.cellMask, np.float64),
        normalThicknessFlux=f(prob.normalThicknessFlux, np.float64), advMaskHighOrder=f(prob.advMaskHighOrder, np.float64),
        advCoefs=f(prob.advCoefs, np.float64), advCoefs3rd=f(prob.advCoefs3rd, np.float64),
    )
    s = _Problem(prob.nEdges, prob.nCells, prob.nVertLevels, prob.nvldim, prob.nAdv)
    for k, v in keep.items():
        setattr(s, k, v.ctypes.data)
    s.coef3rdOrder = prob.coef3rdOrder
    return s, keep


def _out(prob):
    return np.zeros((prob.nvldim, prob.nEdges), order="F")


def original_cpu(prob):
    """refFlx, nested.F90:119-157."""
    s, keep = _problem(prob)
    out = _out(prob)
    _load("libcke_oracle.so").ckeo_original_cpu(C.byref(s), out.ctypes.data)
    return out


def gpu_optimized(prob):
    """nested.F90:606-630."""
    s, keep = _problem(prob)
    out = _out(prob)
    _load("libcke_oracle.so").ckeo_gpu_optimized(C.byref(s), out.ctypes.data)
    return out


def impl1(prob, tracer=None):
    """One iteration of cke_impl1.cpp:11-37.  Returns (highOrderFlx, tracerCur after the in-place masking)."""
    s, keep = _problem(prob)
    tr = np.array(prob.tracerCur if tracer is None else tracer, dtype=np.float64, order="F", copy=True)
    out = _out(prob)
    _load("libcke_oracle.so").ckeo_impl1(C.byref(s), tr.ctypes.data, out.ctypes.data)
    return out, tr


def impl2(prob, tracer=None):
    """One iteration of cke_impl2.cpp:27-68 at pack size 1."""
    s, keep = _problem(prob)
    tr = np.array(prob.tracerCur if tracer is None else tracer, dtype=np.float64, order="F", copy=True)
    out = _out(prob)
    _load("libcke_oracle.so").ckeo_impl2(C.byref(s), tr.ctypes.data, out.ctypes.data)
    return out


def check(flx, ref, prob, err_tol=1.0e-10):
    """The reference's acceptance test nested.F90:267-279: (points failing errTol, worst error)."""
    worst = C.c_double(0.0)
    flx = np.asfortranarray(flx)
    ref = np.asfortranarray(ref)
    bad = _load("libcke_oracle.so").ckeo_check(
        flx.ctypes.data, ref.ctypes.data, prob.nEdges, prob.nVertLevels, prob.nvldim, err_tol, C.byref(worst)
    )
    return int(bad), worst.value


def time_original_cpu(prob, iters):
    """Seconds per iteration of the original CPU form, OpenMP over edges (timing build, -O3)."""
    import time

    s, keep = _problem(prob)
    out = _out(prob)
    L = _load("libcke_oracle_omp.so")
    L.ckeo_original_cpu_omp(C.byref(s), out.ctypes.data, 1)
    t0 = time.perf_counter()
    L.ckeo_original_cpu_omp(C.byref(s), out.ctypes.data, iters)
    return (time.perf_counter() - t0) / iters, out
