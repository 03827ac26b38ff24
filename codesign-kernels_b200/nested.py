"""ctypes binding of include/cke_b200.h (libcke_b200.so) plus a host-side mirror of the reference
mini-app's driver, `program nested` (nested_loops/nested.F90, cited as nested.F90:<line>).

Names and argument meaning follow nested_loops/cke_mod.F90:4-50; arrays are numpy arrays in
Fortran order with the reference's shapes (`tracerCur(nvldim,nCells)` ...), indices 1-based, so
the bytes handed to the library are exactly what nested.F90:395-398 passes.

There is no CPU fallback: a missing library or a box without a B200 raises `CkeError`.
"""
import ctypes as C
import os
import re

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_HEADER = os.path.join(os.path.dirname(_HERE), "include", "cke_b200.h")

COEF3RD_ORDER = 2.14  # nested_vars.F90:32 (a default-real literal there: see NestedProblem.coef3rdOrder)
ERR_TOL = 1.0e-10  # nested_vars.F90:33


class CkeError(RuntimeError):
    pass


def lib_path():
    return os.path.join(_HERE, "lib", "libcke_b200.so")


def declared_symbols():
    """Every function name declared in include/cke_b200.h."""
    text = re.sub(r"/\*.*?\*/", "", open(_HEADER).read(), flags=re.S)
    return sorted(set(re.findall(r"\b((?:cke|kokkos|yakl)_[A-Za-z0-9_]+)\s*\(", text)))


_lib = None


def lib():
    """Loads libcke_b200.so; fails loudly if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise CkeError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` (nvcc, sm_100a)")
    L = C.CDLL(path)
    i32, p, f64 = C.c_int, C.c_void_p, C.c_double
    L.kokkos_init.argtypes, L.kokkos_init.restype = [], None
    L.kokkos_finalize.argtypes, L.kokkos_finalize.restype = [], None
    L.cke_init.argtypes = [i32] * 6 + [p] * 10 + [f64, p]
    L.cke_init.restype = None
    L.cke_get_results.argtypes, L.cke_get_results.restype = [i32, i32, p], None
    L.cke_cleanup.argtypes, L.cke_cleanup.restype = [], None
    L.cke_impl1_run.argtypes, L.cke_impl1_run.restype = [], None
    L.cke_impl2_run.argtypes, L.cke_impl2_run.restype = [], None
    L.yakl_init.argtypes, L.yakl_init.restype = [], None
    L.yakl_init_arrays.argtypes, L.yakl_init_arrays.restype = [i32] * 5 + [f64], None
    for f in (L.yakl_original_loop, L.yakl_gpu_optimized):
        f.argtypes, f.restype = [p] * 3, None
    L.cke_b200_set_abort_on_error.argtypes, L.cke_b200_set_abort_on_error.restype = [i32], None
    L.cke_b200_last_error.argtypes, L.cke_b200_last_error.restype = [], C.c_char_p
    L.cke_b200_launch_count.argtypes, L.cke_b200_launch_count.restype = [], C.c_longlong
    L.cke_b200_time_ms.argtypes, L.cke_b200_time_ms.restype = [i32, i32, C.POINTER(f64)], i32
    L.cke_b200_time_cold_ms.argtypes, L.cke_b200_time_cold_ms.restype = [i32, i32, C.POINTER(f64)], i32
    L.cke_b200_get_tracer.argtypes, L.cke_b200_get_tracer.restype = [p], i32
    # Python callers get exceptions instead of the reference's abort-on-assert
    L.cke_b200_set_abort_on_error(0)
    _lib = L
    return L


def _check(L, what):
    msg = L.cke_b200_last_error()
    if msg:
        raise CkeError(f"{what}: {msg.decode()}")


class NestedProblem:
    """The arrays of `module nested_vars` (nested_vars.F90:35-68), allocated as `alloc_vars`
    does (nested_vars.F90:101-123) and filled the way `program nested` fills them
    (nested.F90:60-106): bottom levels between 3 and nVertLevels with about half at full depth,
    tracer 15*r inside the active range and 0 outside, cellMask 1/0, random 1-based connectivity,
    advCoefs 20*r, advCoefs3rd 21*r, normalThicknessFlux 15*(0.5-r), advMaskHighOrder 1.

    The reference draws from the compiler's `random_number`, whose stream is not portable; here r
    comes from numpy's PCG64 seeded with `seed`.  `packn` is the reference's F90_PACK_SIZE:
    nvldim = ceil(nVertLevels/packn)*packn, the padding levels are zero.
    """

    def __init__(self, nEdges=25600, nCells=2800, nVertLevels=100, nAdv=10, nIters=100, packn=1, seed=0):
        # defaults: nested_loops/nested.nml:2-6
        self.nIters, self.nEdges, self.nCells, self.nVertLevels, self.nAdv = nIters, nEdges, nCells, nVertLevels, nAdv
        self.packn = packn
        self.nvlpk = (nVertLevels + packn - 1) // packn
        self.nvldim = self.nvlpk * packn
        # `coef3rdOrder = 2.14` is a default-real literal assigned to a double (nested_vars.F90:32):
        # the value every loop form uses is float32(2.14) widened, 2.1400001049041748
        self.coef3rdOrder = float(np.float32(COEF3RD_ORDER))
        rng = np.random.default_rng(seed)
        nv = self.nvldim
        self.minLevelCell = np.ones(nCells, dtype=np.int32)
        k = np.rint(rng.random(nCells) * nVertLevels * 2.0).astype(np.int64)
        self.maxLevelCell = np.minimum(np.maximum(3, k), nVertLevels).astype(np.int32)
        lev = np.arange(1, nv + 1)[:, None]
        active = (lev >= self.minLevelCell[None, :]) & (lev <= self.maxLevelCell[None, :])
        self.tracerCur = np.asfortranarray(np.where(active, 15.0 * rng.random((nv, nCells)), 0.0))
        self.cellMask = np.asfortranarray(active.astype(np.float64))
        self.nAdvCellsForEdge = np.full(nEdges, nAdv, dtype=np.int32)
        self.advCellsForEdge = np.asfortranarray((nCells * rng.random((nAdv, nEdges))).astype(np.int32) + 1)
        if nCells:
            np.minimum(self.advCellsForEdge, nCells, out=self.advCellsForEdge)
        self.advCoefs = np.asfortranarray(20.0 * rng.random((nAdv, nEdges)))
        self.advCoefs3rd = np.asfortranarray(21.0 * rng.random((nAdv, nEdges)))
        self.normalThicknessFlux = np.zeros((nv, nEdges), order="F")
        self.normalThicknessFlux[:nVertLevels] = 15.0 * (0.5 - rng.random((nVertLevels, nEdges)))
        self.advMaskHighOrder = np.zeros((nv, nEdges), order="F")
        self.advMaskHighOrder[:nVertLevels] = 1.0
        self.highOrderFlx = np.zeros((nv, nEdges), order="F")

    def points(self):
        return self.nEdges * self.nVertLevels

    def algorithmic_bytes(self):
        """Compulsory DRAM bytes of one iteration of the flux loop: normalThicknessFlux and
        advMaskHighOrder read, highOrderFlx written (8 B each per point), one (cell, coef, coef3rd)
        record of 20 B per (edge, contributing cell), tracerCur read once (it is gathered from L2)."""
        return 24 * self.nEdges * self.nvldim + 20 * int(self.nAdvCellsForEdge.sum()) + 8 * self.nCells * self.nvldim


class NestedLoops:
    """The C++ leg of `program nested` (nested.F90:391-433): kokkos_init, cke_init, cke_implN_run,
    cke_get_results, cke_cleanup, kokkos_finalize -- same names, same order of use."""

    def __init__(self):
        self.L = lib()
        self._prob = None

    def kokkos_init(self):
        self.L.kokkos_init()
        _check(self.L, "kokkos_init")

    def kokkos_finalize(self):
        self.L.kokkos_finalize()

    def cke_init(self, prob, nIters=None):
        a = lambda x, dt: np.asfortranarray(x, dtype=dt)  # noqa: E731
        self._keep = [
            a(prob.nAdvCellsForEdge, np.int32), a(prob.minLevelCell, np.int32), a(prob.maxLevelCell, np.int32),
            a(prob.advCellsForEdge, np.int32), a(prob.tracerCur, np.float64), a(prob.normalThicknessFlux, np.float64),
            a(prob.advMaskHighOrder, np.float64), a(prob.cellMask, np.float64), a(prob.advCoefs, np.float64),
            a(prob.advCoefs3rd, np.float64),
        ]
        ptr = [x.ctypes.data for x in self._keep]
        self.L.cke_init(
            prob.nIters if nIters is None else nIters, prob.nEdges, prob.nCells, prob.nVertLevels, prob.nvldim, prob.nAdv,
            *ptr, prob.coef3rdOrder, prob.highOrderFlx.ctypes.data,
        )
        _check(self.L, "cke_init")
        self._prob = prob

    def cke_impl1_run(self):
        self.L.cke_impl1_run()
        _check(self.L, "cke_impl1_run")

    def cke_impl2_run(self):
        self.L.cke_impl2_run()
        _check(self.L, "cke_impl2_run")

    def cke_get_results(self, highOrderFlx=None):
        prob = self._prob
        if prob is None:
            raise CkeError("cke_get_results before cke_init")
        out = prob.highOrderFlx if highOrderFlx is None else highOrderFlx
        assert out.flags.f_contiguous and out.shape == (prob.nvldim, prob.nEdges) and out.dtype == np.float64
        self.L.cke_get_results(prob.nEdges, prob.nVertLevels, out.ctypes.data)
        _check(self.L, "cke_get_results")
        return out

    def get_tracer(self):
        prob = self._prob
        out = np.zeros((prob.nvldim, prob.nCells), order="F")
        if self.L.cke_b200_get_tracer(out.ctypes.data):
            _check(self.L, "cke_b200_get_tracer")
        return out

    def time_ms(self, impl, iters):
        ms = C.c_double(0.0)
        if self.L.cke_b200_time_ms(impl, iters, C.byref(ms)):
            _check(self.L, "cke_b200_time_ms")
        return ms.value

    def time_cold_ms(self, impl, reps):
        ms = C.c_double(0.0)
        if self.L.cke_b200_time_cold_ms(impl, reps, C.byref(ms)):
            _check(self.L, "cke_b200_time_cold_ms")
        return ms.value

    def launch_count(self):
        return int(self.L.cke_b200_launch_count())

    def cke_cleanup(self):
        self.L.cke_cleanup()
        self._prob = None


class YaklLoops:
    """The YAKL leg of `program nested` as the Fortran side drives it (nested_vars.F90:75-86,138-151 publishes the arrays
    through the bind(C) pointer globals c_*, nested.F90 then calls yakl_init, yakl_init_arrays and a loop form per iteration):
    same symbol names as the reference's nested_yakl_c.cpp:130-306, served by libcke_b200.so."""

    GLOBALS = (("c_maxLevelCell", "maxLevelCell", np.int32), ("c_minLevelCell", "minLevelCell", np.int32),
               ("c_nAdvCellsForEdge", "nAdvCellsForEdge", np.int32), ("c_advCellsForEdge", "advCellsForEdge", np.int32),
               ("c_advCoefs", "advCoefs", np.float64), ("c_advCoefs3rd", "advCoefs3rd", np.float64),
               ("c_cellMask", "cellMask", np.float64), ("c_advMaskHighOrder", "advMaskHighOrder", np.float64))

    def __init__(self, prob):
        self.L = lib()
        self.prob = prob
        self.L.yakl_init()
        _check(self.L, "yakl_init")
        self._keep = {}
        for sym, attr, dt in self.GLOBALS:
            arr = np.asfortranarray(getattr(prob, attr), dtype=dt)
            self._keep[sym] = arr
            C.c_void_p.in_dll(self.L, sym).value = arr.ctypes.data
        self.L.yakl_init_arrays(prob.nvldim, prob.nEdges, prob.nCells, prob.nVertLevels, prob.nAdv, prob.coef3rdOrder)
        _check(self.L, "yakl_init_arrays")

    def _run(self, fn, name, tracer):
        p = self.prob
        tr = np.array(p.tracerCur if tracer is None else tracer, dtype=np.float64, order="F", copy=True)
        ntf = np.array(p.normalThicknessFlux, dtype=np.float64, order="F", copy=True)
        out = np.full((p.nvldim, p.nEdges), np.nan, order="F")
        fn(out.ctypes.data, tr.ctypes.data, ntf.ctypes.data)
        _check(self.L, name)
        assert np.array_equal(tr, p.tracerCur if tracer is None else tracer)  # the host tracer is input only
        return out

    def yakl_original_loop(self, tracer=None):
        return self._run(self.L.yakl_original_loop, "yakl_original_loop", tracer)

    def yakl_gpu_optimized(self, tracer=None):
        return self._run(self.L.yakl_gpu_optimized, "yakl_gpu_optimized", tracer)

    def cleanup(self):
        self.L.cke_cleanup()
