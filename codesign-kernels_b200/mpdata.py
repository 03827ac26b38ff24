"""ctypes binding of include/mpdata_b200.h (libmpdata_b200.so) plus a host-side mirror of the
reference mini-app's driver, `program test_advect`
(mmf-mpdata-tracer/advect_scalar2D_pushncols_openacc.F90, cited as A:<line>).

Arrays are numpy arrays in Fortran order with the reference's shapes (A:14-36; the halo offsets
-2 / -1 of the x index are implicit: column 0 of `f` is x = -2, column 0 of `u`/`w` is x = -1),
so the bytes handed to the library are exactly what A:49-66 passes to the subroutines.

There is no CPU fallback: a missing library or a box without a B200 raises `MpdataError`.
"""
import ctypes as C
import os
import re

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_HEADER = os.path.join(os.path.dirname(_HERE), "include", "mpdata_b200.h")

NSLICES, NZ, NX = 48, 58, 32  # A:6-8 "User changeable"


class MpdataError(RuntimeError):
    pass


def lib_path():
    return os.path.join(_HERE, "lib", "libmpdata_b200.so")


def declared_symbols():
    """Every function name declared in include/mpdata_b200.h."""
    text = re.sub(r"/\*.*?\*/", "", open(_HEADER).read(), flags=re.S)
    return sorted(set(re.findall(r"\b(mpdata_[A-Za-z0-9_]+)\s*\(", text)))


_lib = None


def lib():
    """Loads libmpdata_b200.so; fails loudly if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise MpdataError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` (nvcc, sm_100a)")
    L = C.CDLL(path)
    i32, p = C.c_int, C.c_void_p
    L.mpdata_advect_scalar2D.argtypes, L.mpdata_advect_scalar2D.restype = [i32] * 3 + [p] * 7, i32
    L.mpdata_init.argtypes, L.mpdata_init.restype = [i32] * 3, i32
    L.mpdata_upload.argtypes, L.mpdata_upload.restype = [p] * 7, i32
    L.mpdata_run.argtypes, L.mpdata_run.restype = [], i32
    L.mpdata_download.argtypes, L.mpdata_download.restype = [p, p], i32
    L.mpdata_finalize.argtypes, L.mpdata_finalize.restype = [], i32
    L.mpdata_time_steps.argtypes, L.mpdata_time_steps.restype = [i32, i32, C.POINTER(C.c_float)], i32
    L.mpdata_set_block.argtypes, L.mpdata_set_block.restype = [i32], i32
    L.mpdata_set_chunks.argtypes, L.mpdata_set_chunks.restype = [i32], i32
    L.mpdata_set_form.argtypes, L.mpdata_set_form.restype = [i32, i32], i32
    L.mpdata_march_wins.argtypes, L.mpdata_march_wins.restype = [C.c_longlong, i32], i32
    L.mpdata_get_launch.argtypes, L.mpdata_get_launch.restype = [C.POINTER(i32)] * 3 + [C.POINTER(C.c_longlong)], i32
    L.mpdata_slice_smem_bytes.argtypes, L.mpdata_slice_smem_bytes.restype = [i32, i32], C.c_longlong
    L.mpdata_launch_count.argtypes, L.mpdata_launch_count.restype = [], C.c_longlong
    L.mpdata_last_error.argtypes, L.mpdata_last_error.restype = [], C.c_char_p
    _lib = L
    return L


def _check(rc):
    if rc:
        msg = lib().mpdata_last_error()
        raise MpdataError(f"libmpdata_b200 error {rc}: {msg.decode() if msg else '?'}")


def shapes(nslices=NSLICES, nx=NX, nz=NZ):
    """Array shapes of A:31-36 (the singleton y dimension kept, as in the reference)."""
    nzm = nz - 1
    return {
        "f": (nslices, nx + 6, 1, nzm),
        "u": (nslices, nx + 5, 1, nzm),
        "w": (nslices, nx + 4, 1, nz),
        "rho": (nslices, nzm),
        "rhow": (nslices, nz),
        "adz": (nslices, nzm),
        "flux": (nslices, nz),
    }


class TestAdvect:
    """Mirror of `program test_advect` (A:2-66): init(), the advection call, save()/compare()."""

    __test__ = False  # not a pytest class

    ORDER = ("adz", "f", "u", "w", "rho", "rhow", "flux")  # order of the random_number calls, A:654-660

    def __init__(self, nslices=NSLICES, nx=NX, nz=NZ):
        self.nslices, self.nx, self.nz = nslices, nx, nz
        self.shapes = shapes(nslices, nx, nz)

    def init(self, seed=100, signed_velocity=False):
        """A:645-665: uniform [0,1) fields in the reference's draw order.  The reference seeds the
        compiler's generator (`seed = 100`), which cannot be reproduced; numpy's generator is used.
        `signed_velocity` (not in the reference) shifts u and w to [-0.5,0.5) so that both upwind
        branches and both signs of the limiter run."""
        rng = np.random.default_rng(seed)
        for name in self.ORDER:
            a = rng.random(int(np.prod(self.shapes[name]))).reshape(self.shapes[name], order="F")
            if signed_velocity and name in ("u", "w"):
                a = a - 0.5
            setattr(self, name, np.asfortranarray(a))
        return self

    def arrays(self):
        return [getattr(self, n) for n in ("f", "u", "w", "rho", "rhow", "adz", "flux")]

    def advect_scalar2D(self):
        """The call of A:49 / A:53 with the B200 library behind it: f and flux(:,1:nzm) updated in place."""
        for a in self.arrays():
            assert a.flags.f_contiguous and a.dtype == np.float64
        f, u, w, rho, rhow, adz, flux = self.arrays()
        _check(lib().mpdata_advect_scalar2D(self.nslices, self.nx, self.nz, *[a.ctypes.data for a in (f, u, w, rho, rhow, adz, flux)]))
        return self

    def save(self):
        """A:668-676"""
        self.f_save, self.flux_save = self.f.copy(order="F"), self.flux.copy(order="F")

    def compare(self):
        """A:679-683: relative L1 errors of f and flux against the saved run."""
        rel = lambda a, b: float(np.abs(a - b).sum() / np.abs(b).sum())
        return rel(self.f, self.f_save), rel(self.flux, self.flux_save)


def launch_shape():
    """(grid, block, chunks, shared-memory bytes) the library uses for the sizes of the last mpdata_init."""
    g, b, c, s = C.c_int(), C.c_int(), C.c_int(), C.c_longlong()
    _check(lib().mpdata_get_launch(C.byref(g), C.byref(b), C.byref(c), C.byref(s)))
    return g.value, b.value, c.value, s.value


def set_form(form=0, warps_per_cta=0):
    """Which kernel runs the step: 0 the library's choice, 1 barrier-phased CTA per slice, 2 / 3 marching warp per slice /
    per two slices (reference sizes only).  Results do not depend on it."""
    _check(lib().mpdata_set_form(form, warps_per_cta))


def time_steps(prob, iters=50, flush_l2=False, block=None, chunks=None, form=None, warps_per_cta=0):
    """Device-resident timing of one step at prob's sizes: mean kernel time in ms (CUDA events).
    block / chunks / form: None leaves the library's setting, 0 = automatic."""
    L = lib()
    if form is not None:
        _check(L.mpdata_set_form(form, warps_per_cta))
    if block is not None:
        _check(L.mpdata_set_block(block))
    if chunks is not None:
        _check(L.mpdata_set_chunks(chunks))
    _check(L.mpdata_init(prob.nslices, prob.nx, prob.nz))
    f, u, w, rho, rhow, adz, flux = prob.arrays()
    _check(L.mpdata_upload(*[a.ctypes.data for a in (f, u, w, rho, rhow, adz, flux)]))
    ms = C.c_float()
    _check(L.mpdata_time_steps(iters, int(bool(flush_l2)), C.byref(ms)))
    return ms.value


def algorithmic_bytes(nslices=NSLICES, nx=NX, nz=NZ):
    """HBM bytes one step must move: f, u, w, rho, rhow, adz read once; f (the x = -1..nx+2 columns the
    step changes) and flux(:,1:nzm) written once."""
    nzm = nz - 1
    rd = (nx + 6) * nzm + (nx + 5) * nzm + (nx + 4) * nz + 3 * nzm
    wr = (nx + 4) * nzm + nzm
    return 8 * nslices * (rd + wr)
