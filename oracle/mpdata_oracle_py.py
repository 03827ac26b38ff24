"""TEST INFRASTRUCTURE ONLY -- ctypes view of oracle/mpdata_oracle.c (advect_scalar2D_cpu restated).

Only tests/, __graft_entry__ and bench / profiles cpu_baseline legs may import this module; the
product package never does.  Pinned bit for bit on the reference's source translated to C (oracle/ref_mpdata_py.py).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_libs = {}


def _load(name):
    if name in _libs:
        return _libs[name]
    path = os.path.join(_HERE, name)
    if not os.path.exists(path):
        subprocess.run(["make", "-C", _HERE], check=True, stdout=subprocess.DEVNULL)
    L = C.CDLL(path)
    p, i32 = C.c_void_p, C.c_int
    L.mpdo_advect_scalar2D_cpu.argtypes, L.mpdo_advect_scalar2D_cpu.restype = [i32] * 3 + [p] * 7, i32
    L.mpdo_advect_scalar2D_cpu_blocks.argtypes, L.mpdo_advect_scalar2D_cpu_blocks.restype = [i32] * 3 + [p] * 7 + [i32], i32
    L.mpdo_relative_l1.argtypes, L.mpdo_relative_l1.restype = [p, p, C.c_size_t], C.c_double
    _libs[name] = L
    return L


def advect_scalar2D_cpu(nslices, nx, nz, f, u, w, rho, rhow, adz, flux, omp_block=0, omp=False):
    """In place on f and flux (Fortran-ordered float64 arrays of the reference's shapes).
    omp=True uses the timing build (-O3 -fopenmp: threads take levels inside every loop nest, same bits);
    omp_block > 0 uses the timing build on blocks of that many slices gathered into private copies."""
    arrs = (f, u, w, rho, rhow, adz, flux)
    for a in arrs:
        assert a.flags.f_contiguous and a.dtype == np.float64
    ptrs = [a.ctypes.data for a in arrs]
    if omp:
        rc = _load("libmpdata_oracle_omp.so").mpdo_advect_scalar2D_cpu(nslices, nx, nz, *ptrs)
    elif omp_block:
        rc = _load("libmpdata_oracle_omp.so").mpdo_advect_scalar2D_cpu_blocks(nslices, nx, nz, *ptrs, omp_block)
    else:
        rc = _load("libmpdata_oracle.so").mpdo_advect_scalar2D_cpu(nslices, nx, nz, *ptrs)
    if rc:
        raise ValueError("mpdata oracle: bad sizes")


def relative_l1(a, b):
    a, b = np.ascontiguousarray(a.ravel(order="K")), np.ascontiguousarray(b.ravel(order="K"))
    return _load("libmpdata_oracle.so").mpdo_relative_l1(a.ctypes.data, b.ctypes.data, a.size)
