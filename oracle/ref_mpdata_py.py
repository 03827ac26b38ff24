"""TEST INFRASTRUCTURE ONLY -- ctypes view of oracle/_ref/libmpdata_f90ref.so: the REFERENCE'S OWN Fortran subroutines of
the mmf-mpdata-tracer mini-app (mmf-mpdata-tracer/advect_scalar2D_pushncols_openacc.F90, cited A:<line>)

    advect_scalar2D_cpu        A:477-642   the CPU form the reference's acceptance test compares against
    advect_scalar2D_openacc_1  A:247-474   the 16-kernel OpenACC form (directives are comments to a non-OpenACC compiler)
    advect_scalar2D_openacc_2  A:72-244    the 9-kernel OpenACC form

translated mechanically, statement by statement, by oracle/f90_to_c.py from the source where it lies under /root/reference
(no Fortran compiler exists in this image; nothing of the reference is copied into the repo) and compiled by
`make -C oracle ref` with gcc -O2 -ffp-contract=off.  It pins oracle/mpdata_oracle.c (bit for bit, tests/test_zzz_mpdata.py)
and, through the GPU tests and smoke(), libmpdata_b200.so.  What it is NOT: the output of a Fortran compiler -- the claim is
"the reference's statements, in the reference's order, with C's (= Fortran's) mixed-mode conversions and gfortran's
left-to-right max/min", which is exactly what a hand restatement can get wrong and a compiler cannot.

advect_scalar2D_openacc_2 is built with the translator's --check-bounds: its line A:126 writes www(sl,nxp3,1,nz) for every
slice, one column past www's declared upper bound nxp2; the checked build skips and counts those statements
(violations() below), everything else runs as written.

Only tests/, __graft_entry__.smoke() and tests/golden/make_mpdata_golden.py may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
PATH = os.path.join(_HERE, "_ref", "libmpdata_f90ref.so")
SOURCE = "/root/reference/mmf-mpdata-tracer/advect_scalar2D_pushncols_openacc.F90"
FORMS = ("cpu", "openacc_1", "openacc_2")
_lib = None


def available():
    """True if the translated reference exists (or can be built because /root/reference is here)."""
    if os.path.exists(PATH):
        return True
    if os.path.exists(SOURCE):
        subprocess.run(["make", "-C", _HERE, "ref"], check=False, stdout=subprocess.DEVNULL)
    return os.path.exists(PATH)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libmpdata_f90ref.so is absent and /root/reference is not here to build it")
        L = C.CDLL(PATH)
        for form in FORMS:
            f = getattr(L, f"advect_scalar2d_{form}")
            f.argtypes, f.restype = [C.c_int] * 3 + [C.c_void_p] * 7, C.c_int
        L.advect_scalar2d_openacc_2_violations.argtypes = [C.POINTER(C.c_int)] * 4 + [C.POINTER(C.c_char_p)]
        L.advect_scalar2d_openacc_2_violations.restype = C.c_long
        _lib = L
    return _lib


def advect_scalar2D(form, nslices, nx, nz, f, u, w, rho, rhow, adz, flux):
    """One step of the reference's subroutine `advect_scalar2D_<form>` on Fortran-ordered arrays of the reference's shapes
    (A:31-36); f and flux(:,1:nzm) are updated in place.  Argument order as oracle/mpdata_oracle_py.advect_scalar2D_cpu."""
    assert form in FORMS
    for a in (f, u, w, rho, rhow, adz, flux):
        assert a.flags.f_contiguous and a.dtype == np.float64
    rc = getattr(lib(), f"advect_scalar2d_{form}")(nslices, nz, nx, *[a.ctypes.data for a in (f, u, w, rho, rhow, flux, adz)])
    if rc:
        raise RuntimeError(f"translated advect_scalar2D_{form} returned {rc} (out of memory)")


def advect_scalar2D_cpu(nslices, nx, nz, f, u, w, rho, rhow, adz, flux):
    advect_scalar2D("cpu", nslices, nx, nz, f, u, w, rho, rhow, adz, flux)


def violations():
    """Out-of-bounds statements skipped by the last advect_scalar2D_openacc_2 call:
    (count, source line, array, dimension, subscript, declared upper bound)."""
    a = [C.c_int() for _ in range(4)]
    name = C.c_char_p()
    n = lib().advect_scalar2d_openacc_2_violations(*[C.byref(x) for x in a], C.byref(name))
    return n, a[0].value, (name.value or b"").decode(), a[1].value, a[2].value, a[3].value
