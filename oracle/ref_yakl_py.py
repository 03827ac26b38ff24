"""TEST INFRASTRUCTURE ONLY -- ctypes view of oracle/_ref/libnested_yakl_ref.so: the REFERENCE'S OWN
C++ implementation of the nested_loops flux loop (nested_loops/nested_yakl_c.cpp:142-306 on the
vendored YAKL headers, host back end), compiled by `make -C oracle ref` from the sources where they
lie under /root/reference (never copied).  This is "the reference itself run here": it pins
oracle/cke_oracle.c and, through the GPU tests, libcke_b200.so.

The Fortran driver publishes its arrays through `type(c_ptr), bind(C)` globals
(nested_vars.F90:7,75-86,138-151 -> the c_* symbols at nested_yakl_c.cpp:130-140), then calls
yakl_init (nested.F90 via nested_vars), yakl_init_arrays and the two loop forms; load() does the same.

Only tests/, __graft_entry__.smoke() and profiles/ cpu legs may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
PATH = os.path.join(_HERE, "_ref", "libnested_yakl_ref.so")
_lib = None


def available():
    """True if the reference-built library exists (or can be built because /root/reference is here)."""
    if os.path.exists(PATH):
        return True
    if os.path.exists("/root/reference/nested_loops/nested_yakl_c.cpp"):
        subprocess.run(["make", "-C", _HERE, "ref"], check=False, stdout=subprocess.DEVNULL)
    return os.path.exists(PATH)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libnested_yakl_ref.so is absent and /root/reference is not here to build it")
        L = C.CDLL(PATH)
        L.yakl_init.argtypes, L.yakl_init.restype = [], None
        L.yakl_init_arrays.argtypes = [C.c_int] * 5 + [C.c_double]
        L.yakl_init_arrays.restype = None
        for f in (L.yakl_original_loop, L.yakl_gpu_optimized):
            f.argtypes, f.restype = [C.c_void_p] * 3, None
        L.yakl_init()
        _lib = L
    return _lib


class RefProblem:
    """One `yakl_init_arrays` instance (nested_yakl_c.cpp:149-181).  The library copies the connectivity, coefficients and
    masks at init and takes tracerCur / normalThicknessFlux at every call (nested_yakl_c.cpp:187-188, 264-265)."""

    def __init__(self, prob):
        L = lib()
        f = lambda x, dt: np.asfortranarray(x, dtype=dt)  # noqa: E731
        self.prob = prob
        self.keep = dict(
            c_maxLevelCell=f(prob.maxLevelCell, np.int32), c_minLevelCell=f(prob.minLevelCell, np.int32),
            c_nAdvCellsForEdge=f(prob.nAdvCellsForEdge, np.int32), c_advCellsForEdge=f(prob.advCellsForEdge, np.int32),
            c_advCoefs=f(prob.advCoefs, np.float64), c_advCoefs3rd=f(prob.advCoefs3rd, np.float64),
            c_cellMask=f(prob.cellMask, np.float64), c_advMaskHighOrder=f(prob.advMaskHighOrder, np.float64),
        )
        for name, arr in self.keep.items():
            C.c_void_p.in_dll(L, name).value = arr.ctypes.data
        L.yakl_init_arrays(prob.nvldim, prob.nEdges, prob.nCells, prob.nVertLevels, prob.nAdv, prob.coef3rdOrder)

    def _run(self, fn, tracer):
        p = self.prob
        tr = f64(p.tracerCur if tracer is None else tracer)
        ntf = f64(p.normalThicknessFlux)
        out = np.zeros((p.nvldim, p.nEdges), order="F")
        fn(out.ctypes.data, tr.ctypes.data, ntf.ctypes.data)
        return out

    def original_loop(self, tracer=None):
        """yakl_original_loop, nested_yakl_c.cpp:184-243 (level ranges of the cell)."""
        return self._run(lib().yakl_original_loop, tracer)

    def gpu_optimized(self, tracer=None):
        """yakl_gpu_optimized, nested_yakl_c.cpp:260-306 (cellMask multiplication)."""
        return self._run(lib().yakl_gpu_optimized, tracer)


def f64(x):
    return np.array(x, dtype=np.float64, order="F", copy=True)
