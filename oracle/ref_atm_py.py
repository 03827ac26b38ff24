"""TEST INFRASTRUCTURE ONLY -- ctypes view of oracle/_ref/libatm_f90ref.so: the REFERENCE'S OWN three implementations of the
hot path (atmosphere/biharmonic_wk_kernel.F90, cited F90:<line>)

    module biharmonic_wk_scalar_cpu                   F90:100-202   the CPU form (array-valued functions on one level)
    module biharmonic_wk_scalar_gpu_compiler_inline   F90:209-362   "GPU1": blocks of kchunk = 8 levels down the call stack
    module biharmonic_wk_scalar_gpu_push_loop         F90:369-536   "GPU2": three sweeps over all (element, level) pairs
    module kernel_utils                               F90:48-91     initialize_data / myrandom, save_qtens, compute_l2norm

translated mechanically, statement by statement, by oracle/f90_to_c.py (unit mode) from the source where it lies under
/root/reference -- no Fortran compiler exists in this image, nothing of the reference is copied into the repo -- and compiled by
`make -C oracle ref` with gcc -O2 -ffp-contract=off (OpenACC directives are comments, as they are to gfortran without -fopenacc).

Why this can be trusted although no Fortran compiler produced it: on the reference's own LCG inputs (F90:48-58, 77-91) the
translated GPU1 and GPU2 forms differ from the translated CPU form by relative L2 norms of 1.4477731365593603E-016 and
1.3341964743086566E-016 -- the 17 digits that gfortran 4.9.3, gfortran 7.1.0 and pgfortran 17.7 all printed
(atmosphere/results/atm.*, tests/golden/kat.json).  Those numbers are pure rounding noise, so they come out only if every
operation of all three forms is done in the reference's order with the reference's (single-precision) rrearth.  With the
harness routines of kernel_utils translated too (class Program below), nothing but translated reference code stands between the
memory and those printed numbers.

The translated forms pin oracle/bwk_oracle.c bit for bit on arbitrary inputs (tests/test_zzy_atm_ref.py) and are what the GPU
test there and smoke() compare the CUDA result with.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg
may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
PATH = os.path.join(_HERE, "_ref", "libatm_f90ref.so")
SOURCE = "/root/reference/atmosphere/biharmonic_wk_kernel.F90"
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
            raise RuntimeError("oracle/_ref/libatm_f90ref.so is absent and /root/reference is not here to build it")
        L = C.CDLL(PATH)
        i, p = C.c_int, C.c_void_p
        L.atm_cpu_biharmonic_wk_scalar.argtypes, L.atm_cpu_biharmonic_wk_scalar.restype = [i, i, p, p, p, i, i], i
        L.atm_gpu1_biharmonic_wk_scalar.argtypes, L.atm_gpu1_biharmonic_wk_scalar.restype = [i, i, p, p, p, i, i], i
        L.atm_gpu2_biharmonic_wk_scalar.argtypes, L.atm_gpu2_biharmonic_wk_scalar.restype = [i, i, i, p, p, p, p, i, i], i
        _lib = L
    return _lib


class Program:
    """`program biharmonic_wk_scalar_kernel` (F90:545-580) on translated code only: the module variables of kernel_utils
    (elem, deriv, qtens, qtens_save, grads; F90:29-33) are numpy arrays bound to the translated module, at the reference's
    compile-time sizes np = 4, nlev = 72, qsize = 40 and its default nelemd = nets..nete = 16 (F90:10-17); initialize_data
    (F90:48-58, with myrandom F90:77-91), save_qtens (F90:62-65) and compute_l2norm (F90:69-73) are the translated routines."""

    NELEMD, NLEV, QSIZE = 16, 72, 40

    def __init__(self):
        L = lib()
        L.atm_utils_bind.argtypes, L.atm_utils_bind.restype = [C.c_char_p, C.c_void_p], C.c_int
        L.atm_utils_compute_l2norm.argtypes, L.atm_utils_compute_l2norm.restype = [], C.c_double
        self.elem = np.full((self.NELEMD, 144), np.nan)
        self.deriv = np.full(16, np.nan)
        self.qtens = np.full((4, 4, self.NLEV, self.QSIZE, self.NELEMD), np.nan, order="F")
        self.qtens_save = np.full_like(self.qtens, np.nan)
        self.grads = np.full(2 * self.qtens.size, np.nan)
        for name in ("elem", "deriv", "qtens", "qtens_save", "grads"):  # allocate_data, F90:37-44: the memory is numpy's
            assert L.atm_utils_bind(name.encode(), getattr(self, name).ctypes.data) == 0, name

    def initialize_data(self):
        assert lib().atm_utils_initialize_data() == 0

    def save_qtens(self):
        assert lib().atm_utils_save_qtens() == 0

    def compute_l2norm(self):
        return lib().atm_utils_compute_l2norm()

    def run(self):
        """F90:551-578 with the commented GPU1 block (F90:562-569) restored: the three L2-norm-bearing lines the reference's
        results files hold, as (label, value)."""
        out = []
        self.initialize_data()                                          # F90:555
        biharmonic_wk_scalar_cpu(self.elem, self.qtens, self.deriv)     # F90:557  call bws_cpu(elem,qtens,deriv,nets,nete)
        self.save_qtens()                                               # F90:559
        self.initialize_data()                                          # F90:562
        biharmonic_wk_scalar_gpu1(self.elem, self.qtens, self.deriv)    # F90:565
        out.append(("GPU1 L2 norm", self.compute_l2norm()))             # F90:569
        self.initialize_data()                                          # F90:571
        biharmonic_wk_scalar_gpu2(self.elem, self.qtens, self.deriv)    # F90:574
        out.append(("GPU2 L2 norm", self.compute_l2norm()))             # F90:578
        return out


def _check(elem, qtens, dvv):
    assert qtens.flags.f_contiguous and qtens.dtype == np.float64 and qtens.shape[:2] == (4, 4)
    assert elem.dtype == np.float64 and elem.flags.c_contiguous and elem.size == 144 * qtens.shape[4]
    assert dvv.dtype == np.float64 and dvv.size == 16
    return qtens.shape[2], qtens.shape[3], qtens.shape[4]


def biharmonic_wk_scalar_cpu(elem, qtens, dvv, nets=None, nete=None):
    """subroutine biharmonic_wk_scalar of the CPU module (F90:186-200) in place on qtens (4,4,nlev,qsize,nelemd), Fortran
    order; elem is [nelemd][144] (element_t, F90:23-27), dvv the 16 doubles of deriv%Dvv.  Its dummy is
    qtens(np,np,nlev,qsize,nets:nete): the first slab it is handed is element `nets`."""
    nlev, qsize, nelemd = _check(elem, qtens, dvv)
    nets, nete = nets or 1, nete or nelemd
    lib().atm_cpu_biharmonic_wk_scalar(nlev, qsize, elem.ctypes.data, qtens.ctypes.data + 8 * 16 * nlev * qsize * (nets - 1), dvv.ctypes.data, nets, nete)
    return qtens


def biharmonic_wk_scalar_gpu1(elem, qtens, dvv):
    """GPU1 module, F90:317-360 (levels beyond nlev/8*8 are left untouched, as in the source)."""
    nlev, qsize, nelemd = _check(elem, qtens, dvv)
    lib().atm_gpu1_biharmonic_wk_scalar(nlev, qsize, elem.ctypes.data, qtens.ctypes.data, dvv.ctypes.data, 1, nelemd)
    return qtens


def biharmonic_wk_scalar_gpu2(elem, qtens, dvv, nets=None, nete=None):
    """GPU2 module, F90:525-534, with its grads(np,np,2,nlev,qsize,nelemd) scratch (NaN-filled here: it is write-before-read)."""
    nlev, qsize, nelemd = _check(elem, qtens, dvv)
    grads = np.full(2 * qtens.size, np.nan)
    lib().atm_gpu2_biharmonic_wk_scalar(nlev, qsize, nelemd, elem.ctypes.data, qtens.ctypes.data, grads.ctypes.data, dvv.ctypes.data, nets or 1, nete or nelemd)
    return qtens
