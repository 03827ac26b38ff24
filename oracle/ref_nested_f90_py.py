"""TEST INFRASTRUCTURE ONLY -- ctypes view of oracle/_ref/libnested_f90ref.so: the FORTRAN loop forms of the nested_loops
mini-app, internal subroutines of `program nested` (nested_loops/nested.F90)

    run_original_cpu_directive    nested.F90:495-564   per edge, every contributing cell adds on its own level range
    run_gpu_optimized_directive   nested.F90:593-630   collapsed (edge, level) loop, tracer * cellMask

translated mechanically by oracle/f90_to_c.py (state mode: the routines take no arguments and work on the allocatable arrays
of `module nested_vars`, nested_vars.F90:35-68, shaped as its `alloc_vars` allocates them, nested_vars.F90:110-130) from the
sources where they lie, plain CPU build (no -D), and compiled by `make -C oracle ref`.

What it is for: this is the one place where the translator can be held against COMPILED reference code -- the reference's
own C++ forms of the same loops (nested_yakl_c.cpp:184-306, built by g++ into oracle/_ref/libnested_yakl_ref.so).
tests/test_zz_nested.py requires the two to agree bit for bit.  Only tests/ may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
PATH = os.path.join(_HERE, "_ref", "libnested_f90ref.so")
SOURCE = "/root/reference/nested_loops/nested.F90"
_lib = None


def available():
    if os.path.exists(PATH):
        return True
    if os.path.exists(SOURCE):
        subprocess.run(["make", "-C", _HERE, "ref"], check=False, stdout=subprocess.DEVNULL)
    return os.path.exists(PATH)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libnested_f90ref.so is absent and /root/reference is not here to build it")
        L = C.CDLL(PATH)
        L.nested_f90_set_int.argtypes, L.nested_f90_set_int.restype = [C.c_char_p, C.c_int], C.c_int
        L.nested_f90_bind.argtypes, L.nested_f90_bind.restype = [C.c_char_p, C.c_void_p], C.c_int
        for r in ("run_original_cpu_directive", "run_gpu_optimized_directive"):
            getattr(L, "nested_f90_" + r).argtypes, getattr(L, "nested_f90_" + r).restype = [], C.c_int
        _lib = L
    return _lib


def run(prob, form, tracer=None):
    """highOrderFlx(nvldim, nEdges) of `form` ("original" or "gpu_optimized") on a codesign_kernels_b200.nested.NestedProblem:
    the module's integers are set as alloc_vars / the namelist would (nested_vars.F90:104-108), its arrays are bound to copies
    of the problem's, the routine runs.  coef3rdOrder is the module's own parameter (nested_vars.F90:32, a default-real 2.14)."""
    L = lib()
    for name, v in dict(nedges=prob.nEdges, ncells=prob.nCells, nvertlevels=prob.nVertLevels, nadv=prob.nAdv, nvldim=prob.nvldim).items():
        assert L.nested_f90_set_int(name.encode(), v) == 0, name
    f = lambda x, dt: np.array(x, dtype=dt, order="F", copy=True)  # noqa: E731
    nv = prob.nVertLevels
    keep = dict(
        nadvcellsforedge=f(prob.nAdvCellsForEdge, np.int32), minlevelcell=f(prob.minLevelCell, np.int32), maxlevelcell=f(prob.maxLevelCell, np.int32),
        advcellsforedge=f(prob.advCellsForEdge, np.int32), tracercur=f(prob.tracerCur if tracer is None else tracer, np.float64),
        cellmask=f(prob.cellMask, np.float64), normalthicknessflux=f(prob.normalThicknessFlux, np.float64),
        advmaskhighorder=f(prob.advMaskHighOrder, np.float64), advcoefs=f(prob.advCoefs, np.float64), advcoefs3rd=f(prob.advCoefs3rd, np.float64),
        highorderflx=np.zeros((prob.nvldim, prob.nEdges), order="F"),
        wgttmp=np.full(nv, np.nan), sgntmp=np.full(nv, np.nan), flxtmp=np.full(nv, np.nan))  # scratch vectors, undefined on entry
    for name, a in keep.items():
        assert L.nested_f90_bind(name.encode(), a.ctypes.data) == 0, name
    routine = {"original": "run_original_cpu_directive", "gpu_optimized": "run_gpu_optimized_directive"}[form]
    assert getattr(L, "nested_f90_" + routine)() == 0
    return keep["highorderflx"]
