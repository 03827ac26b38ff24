"""TEST INFRASTRUCTURE ONLY -- ctypes view of the C oracle (oracle/bwk_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


def build(quiet=True):
    subprocess.run(["make", "-C", _HERE], check=True, stdout=subprocess.DEVNULL if quiet else None)


def _load(name):
    path = os.path.join(_HERE, name)
    if not os.path.exists(path):
        build()
    L = C.CDLL(path)
    i64, i32, p = C.c_int64, C.c_int, C.c_void_p
    L.bwko_rrearth.restype = C.c_double
    L.bwko_initialize_data.argtypes = [i32, i32, i32, p, p, p]
    L.bwko_initialize_subset.argtypes = [i64, i32, i32, i64, i64, p, p, p]
    L.bwko_biharmonic_wk_scalar_cpu.argtypes = [p, p, p, i32, i32, i32, i32]
    L.bwko_biharmonic_wk_scalar_cpu_omp.argtypes = [p, p, p, i32, i32, i32, i32]
    L.bwko_biharmonic_wk_scalar_gpu1.argtypes = [p, p, p, i32, i32, i32, i32]
    L.bwko_biharmonic_wk_scalar_gpu2.argtypes = [p, p, p, p, i32, i32, i32, i32, i32]
    L.bwko_compute_l2norm.argtypes = [p, p, i64]
    L.bwko_compute_l2norm.restype = C.c_double
    L.bwko_max_rel.argtypes = [p, p, i64]
    L.bwko_max_rel.restype = C.c_double
    L.bwko_cubed_sphere_gdof.argtypes = [i32, p]
    L.bwko_dss_create.argtypes = [i64, p]
    L.bwko_dss_create.restype = p
    L.bwko_dss_destroy.argtypes = [p]
    L.bwko_dss_unique_nodes.argtypes = [p]
    L.bwko_dss_unique_nodes.restype = i64
    L.bwko_dss_multiplicity_hist.argtypes = [p, p]
    L.bwko_dss_rspheremp.argtypes = [p, p, p]
    L.bwko_dss_apply.argtypes = [p, p, i32, i32]
    L.bwko_biharmonic_full.argtypes = [p, p, p, p, p, i32, i32, i32]
    L.bwko_omp_max_threads.restype = i32
    L.bwko_advance_hypervis_scalar.argtypes = [p, p, p, p, p, p, p, i32, i32, C.c_double, C.c_double, i32, i32]
    L.bwko_limiter2d_zero.argtypes = [p, i64]
    L.bwko_biharmonic_wk_dp3d.argtypes = [p, p, p, p, p, p, p, p, i32, i32, i32]
    L.bwko_vlaplace_sphere_wk.argtypes = [p, p, p, p, i32, i64, i32, i32]
    return L


_libs = {}


def lib(omp=False):
    name = "libbwk_oracle_omp.so" if omp else "libbwk_oracle.so"
    if name not in _libs:
        _libs[name] = _load(name)
    return _libs[name]


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def rrearth():
    return lib().bwko_rrearth()


def initialize_data(nelemd, nlev, qsize):
    """F90:48-58 -> (dvv[16], elem[nelemd,144], qtens (4,4,nlev,qsize,nelemd) Fortran order)."""
    dvv = np.empty(16)
    elem = np.empty((nelemd, 144))
    qtens = np.empty((4, 4, nlev, qsize, nelemd), order="F")
    lib().bwko_initialize_data(nelemd, nlev, qsize, _p(dvv), _p(elem), _p(qtens))
    return dvv, elem, qtens


def initialize_subset(nelemd_total, nlev, qsize, first, count):
    dvv = np.empty(16)
    elem = np.empty((count, 144))
    qtens = np.empty((4, 4, nlev, qsize, count), order="F")
    lib().bwko_initialize_subset(nelemd_total, nlev, qsize, first, count, _p(dvv), _p(elem), _p(qtens))
    return dvv, elem, qtens


def initialize_data_parallel(nelemd, nlev, qsize, threads=None, qtens=None):
    """initialize_data (F90:48-58) through the closed form of the LCG (bwko_initialize_subset), element blocks filled by
    host threads (ctypes drops the GIL): the 32 GB ne120 field in seconds instead of a minute.  Same values as
    initialize_data (tests/test_oracle_kat.py)."""
    from concurrent.futures import ThreadPoolExecutor

    threads = threads or max(1, len(os.sched_getaffinity(0)))
    dvv = np.empty(16)
    elem = np.empty((nelemd, 144))
    if qtens is None:
        qtens = np.empty((4, 4, nlev, qsize, nelemd), order="F")
    es = 16 * nlev * qsize
    flat = qtens.reshape(-1, order="F")  # a view: qtens is Fortran-contiguous
    assert flat.base is not None and flat.ctypes.data == qtens.ctypes.data
    step = max(1, -(-nelemd // (threads * 4)))
    L = lib()

    def fill(f):
        c = min(step, nelemd - f)
        d = np.empty(16)
        L.bwko_initialize_subset(nelemd, nlev, qsize, f, c, _p(d), C.c_void_p(elem.ctypes.data + 8 * 144 * f), C.c_void_p(flat.ctypes.data + 8 * es * f))
        return d

    with ThreadPoolExecutor(threads) as ex:
        ds = list(ex.map(fill, range(0, nelemd, step)))
    dvv[:] = ds[0]
    return dvv, elem, qtens


def _dims(qtens):
    return qtens.shape[2], qtens.shape[3], qtens.shape[4]


def biharmonic_wk_scalar_cpu(elem, qtens, dvv, nets=None, nete=None, omp=False):
    """F90:186-200 in place on qtens (Fortran-ordered (4,4,nlev,qsize,nelemd))."""
    nlev, qsize, nelemd = _dims(qtens)
    f = lib(omp).bwko_biharmonic_wk_scalar_cpu_omp if omp else lib().bwko_biharmonic_wk_scalar_cpu
    f(_p(elem), _p(qtens), _p(dvv), nlev, qsize, 1 if nets is None else nets, nelemd if nete is None else nete)
    return qtens


def biharmonic_wk_scalar_gpu1(elem, qtens, dvv):
    nlev, qsize, nelemd = _dims(qtens)
    lib().bwko_biharmonic_wk_scalar_gpu1(_p(elem), _p(qtens), _p(dvv), nlev, qsize, 1, nelemd)
    return qtens


def biharmonic_wk_scalar_gpu2(elem, qtens, dvv):
    nlev, qsize, nelemd = _dims(qtens)
    grads = np.empty(2 * qtens.size)
    lib().bwko_biharmonic_wk_scalar_gpu2(_p(elem), _p(qtens), _p(grads), _p(dvv), nlev, qsize, nelemd, 1, nelemd)
    return qtens


def compute_l2norm(q, qs):
    return lib().bwko_compute_l2norm(_p(q), _p(qs), q.size)


def max_rel(a, b):
    return lib().bwko_max_rel(_p(a), _p(b), a.size)


def cubed_sphere_gdof(ne):
    g = np.empty(16 * 6 * ne * ne, dtype=np.int64)
    lib().bwko_cubed_sphere_gdof(ne, _p(g))
    return g


class Dss:
    def __init__(self, nelem, gdof):
        self.nelem = nelem
        self._g = np.ascontiguousarray(gdof, dtype=np.int64)
        self._h = lib().bwko_dss_create(nelem, _p(self._g))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().bwko_dss_destroy(self._h)
            self._h = None

    def unique_nodes(self):
        return lib().bwko_dss_unique_nodes(self._h)

    def multiplicity_hist(self):
        m = np.zeros(16, dtype=np.int64)
        lib().bwko_dss_multiplicity_hist(self._h, _p(m))
        return m

    def rspheremp(self, elem):
        r = np.empty((4, 4, self.nelem), order="F")
        lib().bwko_dss_rspheremp(self._h, _p(elem), _p(r))
        return r

    def apply(self, qtens):
        nlev, qsize, _ = _dims(qtens)
        lib().bwko_dss_apply(self._h, _p(qtens), nlev, qsize)
        return qtens

    def biharmonic_full(self, elem, rsph, qtens, dvv, omp=False):
        nlev, qsize, _ = _dims(qtens)
        L = lib(omp)
        # the handle belongs to the parity library; the OpenMP build needs its own
        if omp:
            h = L.bwko_dss_create(self.nelem, _p(self._g))
            L.bwko_biharmonic_full(h, _p(elem), _p(rsph), _p(qtens), _p(dvv), nlev, qsize, 1)
            L.bwko_dss_destroy(h)
        else:
            L.bwko_biharmonic_full(self._h, _p(elem), _p(rsph), _p(qtens), _p(dvv), nlev, qsize, 0)
        return qtens

    def advance_hypervis_scalar(self, elem, rsph, qdp, dp, dpdiss, dvv, dt2, nu_q, subcycles=1, limiter=True):
        """upstream HOMME advance_hypervis_scalar (parity unpinned, see bwk_oracle.c); qdp updated in place"""
        nlev, qsize, _ = _dims(qdp)
        lib().bwko_advance_hypervis_scalar(self._h, _p(elem), _p(rsph), _p(qdp), _p(dp), _p(dpdiss), _p(dvv), nlev, qsize,
                                           float(dt2), float(nu_q), int(subcycles), 1 if limiter else 0)
        return qdp

    def biharmonic_wk_dp3d(self, elem, v2c, rsph, dptens, ttens, vtens, dvv, var_coef1=True, undamp=True):
        """upstream HOMME biharmonic_wk_dp3d with the Cartesian vlaplace_sphere_wk (parity unpinned, see bwk_oracle.c);
        dptens, ttens (4,4,nlev,nelem) and vtens (4,4,2,nlev,nelem) Fortran order, updated in place"""
        nlev = ttens.shape[2]
        assert vtens.shape == (4, 4, 2, nlev, self.nelem) and dptens.shape == ttens.shape == (4, 4, nlev, self.nelem)
        lib().bwko_biharmonic_wk_dp3d(self._h, _p(elem), _p(v2c), _p(rsph), _p(dptens), _p(ttens), _p(vtens), _p(dvv), nlev,
                                      1 if var_coef1 else 0, 1 if undamp else 0)
        return dptens, ttens, vtens


def vlaplace_sphere_wk(elem, v2c, v, dvv, var_coef=True, undamp=True):
    """one Cartesian-form vlaplace_sphere_wk per (level, element), in place on v (4,4,2,nlev,nelem)"""
    lib().bwko_vlaplace_sphere_wk(_p(elem), _p(v2c), _p(v), _p(dvv), v.shape[3], v.shape[4], 1 if var_coef else 0, 1 if undamp else 0)
    return v


def limiter2d_zero(q):
    lib().bwko_limiter2d_zero(_p(q), q.size // 16)
    return q
