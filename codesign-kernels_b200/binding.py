"""ctypes binding of include/bwk_b200.h plus the host-side mirror of the reference's
`module kernel_utils` / `biharmonic_wk_scalar` interface (atmosphere/biharmonic_wk_kernel.F90,
cited as F90:<line>).  No torch here: plain pointers and sizes, as a Fortran caller would pass.
"""
import ctypes as C
import os
import re

import numpy as np

NP = 4
_HERE = os.path.dirname(os.path.abspath(__file__))
_HEADER = os.path.join(os.path.dirname(_HERE), "include", "bwk_b200.h")

# type element_t (F90:23-27) and derivative_t (F90:19-21) in C index order
ELEMENT_DTYPE = np.dtype(
    [("Dinv", "<f8", (2, 2, NP, NP)), ("spheremp", "<f8", (NP, NP)), ("tensorVisc", "<f8", (2, 2, NP, NP))]
)
DERIVATIVE_DTYPE = np.dtype([("Dvv", "<f8", (NP, NP))])
assert ELEMENT_DTYPE.itemsize == 144 * 8


class BwkError(RuntimeError):
    pass


def lib_path():
    return os.path.join(_HERE, "lib", "libbwk_b200.so")


def declared_symbols():
    """Every function name declared in include/bwk_b200.h."""
    text = open(_HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bwk_[a-z0-9_]+)\s*\(", text)))


_lib = None


def lib():
    """Loads libbwk_b200.so; fails loudly if it has not been built (no fallback of any kind)."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise BwkError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` (nvcc, sm_100a)")
    L = C.CDLL(path, mode=C.RTLD_GLOBAL)
    i64, i32, p, dp = C.c_int64, C.c_int, C.c_void_p, C.POINTER(C.c_double)
    sig = {
        "bwk_init": [i32, i32, i32],
        "bwk_finalize": [],
        "bwk_upload": [p, p, p],
        "bwk_run": [i32, i32],
        "bwk_download": [p],
        "bwk_sync": [],
        "bwk_biharmonic_wk_scalar": [p, p, p, p, C.POINTER(i32), C.POINTER(i32)],
        "bwk_biharmonic_wk_scalar_nets": [p, p, p, C.POINTER(i32), C.POINTER(i32)],
        "bwk_biharmonic_wk_scalar_full": [p, p, p],
        "bwk_biharmonic_wk_scalar_noalloc": [p, p, p, i32, i32, i32, C.c_uint64],
        "bwk_host_alloc": [C.POINTER(p), C.c_uint64],
        "bwk_host_free": [p],
        "bwk_host_register": [p, C.c_uint64],
        "bwk_host_unregister": [p],
        "bwk_set_variant": [i32],
        "bwk_get_stream": [C.POINTER(C.c_uint64)],
        "bwk_get_device_qtens": [C.POINTER(C.c_uint64)],
        "bwk_event_record": [i32],
        "bwk_event_elapsed_ms": [i32, i32, C.POINTER(C.c_float)],
        "bwk_initialize_data_device": [i64, i64, i32],
        "bwk_save_qtens": [],
        "bwk_compute_l2norm": [dp],
        "bwk_checksum": [dp, dp],
        "bwk_checksum_bits": [i64, C.POINTER(C.c_uint64)],
        "bwk_set_phase_event": [i32],
        "bwk_periodicity_defect": [i64, dp],
        "bwk_download_elements": [i64, i64, p],
        "bwk_upload_elements": [i64, i64, p],
        "bwk_set_topology_gdof": [i64, p, i64],
        "bwk_set_topology_cubed_sphere": [i32, i64],
        "bwk_cubed_sphere_gdof": [i32, p],
        "bwk_set_rspheremp": [p],
        "bwk_get_rspheremp": [p],
        "bwk_run_biharmonic": [],
        "bwk_upload_hypervis_fields": [p, p],
        "bwk_run_advance_hypervis": [C.c_double, C.c_double, i32, i32],
        "bwk_advance_hypervis_scalar": [p, p, p, p, p, C.c_double, C.c_double, i32, i32],
        "bwk_upload_vec_sphere2cart": [p],
        "bwk_run_biharmonic_dp3d": [i32, i32],
        "bwk_biharmonic_wk_dp3d": [p, p, p, p, p, p, i32, i32],
        "bwk_run_biharmonic_phase": [i32],
        "bwk_comm_unique_id": [p],
        "bwk_comm_init": [i32, i32, p],
        "bwk_comm_finalize": [],
        "bwk_comm_barrier": [],
        "bwk_measure_fp64_peak": [C.POINTER(C.c_double)],
        "bwk_partition": [i64, i32, i32, C.POINTER(i64), C.POINTER(i64)],
        "bwk_plan_apply_host": [p, p, p],
        "bwk_plan_send_list": [p, i32, p, p],
        "bwk_plan_recv_list": [p, i32, p, p],
        "bwk_plan_strip_list": [p, i32, p, p, p],
    }
    for name, args in sig.items():
        f = getattr(L, name)
        f.argtypes = args
        f.restype = i32
    L.bwk_last_error.restype = C.c_char_p
    L.bwk_last_error.argtypes = []
    L.bwk_launch_count.restype = i64
    L.bwk_launch_count.argtypes = []
    L.bwk_plan_create.restype = p
    L.bwk_plan_create.argtypes = [i64, p, i32, i32]
    L.bwk_plan_destroy.restype = None
    L.bwk_plan_destroy.argtypes = [p]
    for name in ("bwk_plan_unique_nodes", "bwk_plan_local_first", "bwk_plan_local_count", "bwk_plan_boundary_elems", "bwk_plan_dependent_elems"):
        getattr(L, name).restype = i64
        getattr(L, name).argtypes = [p]
    for name in ("bwk_plan_recv_rows", "bwk_plan_send_rows"):
        getattr(L, name).restype = i64
        getattr(L, name).argtypes = [p, i32]
    L.bwk_plan_strip_count.restype = i64
    L.bwk_plan_strip_count.argtypes = [p, i32]
    L.bwk_plan_strip_groups.restype = i64
    L.bwk_plan_strip_groups.argtypes = [p, i32, i32, p, p]
    _lib = L
    return L


def _ck(rc):
    if rc != 0:
        raise BwkError(lib().bwk_last_error().decode())


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


NOALLOC_MAX_NELEMD = 8192  # BWK_NOALLOC_MAX_NELEMD, include/bwk_b200.h


def biharmonic_wk_scalar_noalloc(d_elem, d_qtens, deriv, nelemd, nlev, qsize, stream=0):
    """Allocation-free single Laplacian on DEVICE pointers (integers); deriv is a host DERIVATIVE_DTYPE array.

    No counterpart in the reference (added for callers that cannot cudaMalloc); needs no bwk_init; asynchronous on `stream`.
    """
    _ck(lib().bwk_biharmonic_wk_scalar_noalloc(C.c_void_p(d_elem), C.c_void_p(d_qtens), _ptr(deriv), nelemd, nlev, qsize, stream))


def cubed_sphere_gdof(ne):
    """gdof[16*e + i + 4*j] of the synthetic cubed sphere; host only."""
    g = np.empty(16 * 6 * ne * ne, dtype=np.int64)
    _ck(lib().bwk_cubed_sphere_gdof(ne, _ptr(g)))
    return g


def partition(nelem_global, rank, nranks):
    f, c = C.c_int64(), C.c_int64()
    _ck(lib().bwk_partition(nelem_global, rank, nranks, C.byref(f), C.byref(c)))
    return f.value, c.value


class Plan:
    """Host-only view of the DSS plan of one rank (no GPU needed)."""

    def __init__(self, nelem_global, gdof, rank=0, nranks=1):
        self.nranks = nranks
        self.rank = rank
        self._gdof = np.ascontiguousarray(gdof, dtype=np.int64)
        self._h = lib().bwk_plan_create(nelem_global, _ptr(self._gdof), rank, nranks)
        if not self._h:
            raise BwkError(lib().bwk_last_error().decode())

    def __del__(self):
        if getattr(self, "_h", None):
            lib().bwk_plan_destroy(self._h)
            self._h = None

    unique_nodes = property(lambda s: lib().bwk_plan_unique_nodes(s._h))
    local_first = property(lambda s: lib().bwk_plan_local_first(s._h))
    local_count = property(lambda s: lib().bwk_plan_local_count(s._h))
    boundary_elems = property(lambda s: lib().bwk_plan_boundary_elems(s._h))
    dependent_elems = property(lambda s: lib().bwk_plan_dependent_elems(s._h))

    def _list(self, peer, which):
        n = getattr(lib(), f"bwk_plan_{which}_rows")(self._h, peer)
        e = np.empty(n, dtype=np.int64)
        r = np.empty(n, dtype=np.int32)
        if n:
            _ck(getattr(lib(), f"bwk_plan_{which}_list")(self._h, peer, _ptr(e), _ptr(r)))
        return e, r

    def send_list(self, peer):
        return self._list(peer, "send")

    def recv_list(self, peer):
        return self._list(peer, "recv")

    def strips(self, which=0):
        """(first local element, count) of every sweep-2 strip; which 0 all, 1 dependent, 2 independent."""
        n = lib().bwk_plan_strip_count(self._h, which)
        f = np.empty(n, dtype=np.int32)
        c = np.empty(n, dtype=np.int32)
        r = np.empty(n, dtype=np.int32)
        if n:
            _ck(lib().bwk_plan_strip_list(self._h, which, _ptr(f), _ptr(c), _ptr(r)))
        self.strip_regular = r
        return f, c

    def strip_groups(self, which=0, max_members=8):
        """(first local element, chained full strips) per warp-strip work unit; 0 strips = left to the round-1 strip kernel."""
        cap = lib().bwk_plan_strip_count(self._h, which)
        f = np.empty(cap, dtype=np.int32)
        k = np.empty(cap, dtype=np.int32)
        n = lib().bwk_plan_strip_groups(self._h, which, max_members, _ptr(f), _ptr(k))
        if n < 0:
            raise ValueError("which must be 0, 1 or 2")
        return f[:n].copy(), k[:n].copy()

    def apply_host(self, field_global):
        f = np.ascontiguousarray(field_global, dtype=np.float64)
        out = np.empty(16 * self.local_count, dtype=np.float64)
        _ck(lib().bwk_plan_apply_host(self._h, _ptr(f), _ptr(out)))
        return out


class KernelUtils:
    """Mirror of `module kernel_utils` + the driver program (F90:7-93, 545-580) for one process.

    Attributes keep the reference's names: np, nlev, qsize, nelemd, nets, nete, elem, deriv,
    qtens, grads.  `qtens` has shape (np,np,nlev,qsize,nelemd) in Fortran order.
    """

    def __init__(self, nelemd=16, nlev=72, qsize=40):  # defaults F90:11-17
        self.np, self.nlev, self.qsize, self.nelemd = NP, nlev, qsize, nelemd
        self.nets, self.nete = 1, nelemd
        self.elem = self.deriv = self.qtens = self.grads = None
        self._pinned = []
        self._open = False

    # -- allocate_data (F90:37-44) + `!$acc enter data pcreate` (F90:553)
    def allocate_data(self, host_arrays=True, pinned=False):
        _ck(lib().bwk_init(self.nelemd, self.nlev, self.qsize))
        self._open = True
        if host_arrays:
            self.elem = np.zeros(self.nelemd, dtype=ELEMENT_DTYPE)
            self.deriv = np.zeros((), dtype=DERIVATIVE_DTYPE)
            shape = (NP, NP, self.nlev, self.qsize, self.nelemd)
            if pinned:
                self.qtens = self._pinned_array(shape)
            else:
                self.qtens = np.zeros(shape, dtype=np.float64, order="F")
            self.grads = None  # F90:42; the fused kernel needs no gradient scratch
        return self

    def _pinned_array(self, shape):
        n = int(np.prod(shape))
        p = C.c_void_p()
        _ck(lib().bwk_host_alloc(C.byref(p), n * 8))
        self._pinned.append(p)
        buf = (C.c_double * n).from_address(p.value)
        return np.frombuffer(buf, dtype=np.float64).reshape(shape, order="F")

    def finalize(self):
        if self._open:
            self.qtens = None
            for p in self._pinned:
                lib().bwk_host_free(p)
            self._pinned = []
            _ck(lib().bwk_finalize())
            self._open = False

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.finalize()

    # -- initialize_data (F90:48-58): host arrays from the reference LCG (myrandom F90:77-91)
    def initialize_data(self):
        state = 11
        def myrandom(n):
            nonlocal state
            out = np.empty(n, dtype=np.float64)
            for t in range(n):
                state = (1301 * state + 97) % (1024 * 128)
                out[t] = state / float(1024 * 128)
            return out
        # vectorised closed form: draw d (0-based) = table[d mod 131072]
        table = myrandom(131072)
        self.deriv["Dvv"] = table[:16].reshape(NP, NP)
        d = 16 + 144 * np.arange(self.nelemd, dtype=np.int64)[:, None] + np.arange(144, dtype=np.int64)[None, :]
        self.elem.view(np.float64).reshape(self.nelemd, 144)[...] = table[d % 131072]
        n = self.qtens.size
        q0 = 16 + 144 * self.nelemd
        flat = self.qtens.reshape(-1, order="F")
        step = 1 << 24
        for a in range(0, n, step):
            b = min(n, a + step)
            flat[a:b] = table[(q0 + np.arange(a, b, dtype=np.int64)) % 131072]
        return self

    # -- device-side twin of initialize_data for the big configurations
    def initialize_data_device(self, nelemd_global=None, elem_first=0, seed_mode=0):
        _ck(lib().bwk_initialize_data_device(self.nelemd if nelemd_global is None else nelemd_global, elem_first, seed_mode))

    # -- `!$acc update device(qtens,elem,deriv)` F90:573
    def update_device(self, elem=True, qtens=True, deriv=True):
        _ck(lib().bwk_upload(_ptr(self.elem) if elem else None, _ptr(self.qtens) if qtens else None, _ptr(self.deriv) if deriv else None))

    # -- `!$acc update host(qtens)` F90:575
    def update_host(self):
        _ck(lib().bwk_download(_ptr(self.qtens)))

    # -- device-resident call, F90:574
    def run_device(self, nets=None, nete=None):
        _ck(lib().bwk_run(self.nets if nets is None else nets, self.nete if nete is None else nete))

    def run_biharmonic_device(self):
        _ck(lib().bwk_run_biharmonic())

    def run_biharmonic_phase(self, phase):
        _ck(lib().bwk_run_biharmonic_phase(phase))

    def sync(self):
        _ck(lib().bwk_sync())

    def measure_fp64_peak(self):
        """measured DFMA rate of this GPU, TFLOP/s"""
        v = C.c_double()
        _ck(lib().bwk_measure_fp64_peak(C.byref(v)))
        return v.value

    def comm_barrier(self):
        """device-side barrier across ranks on the library's stream (no-op on one rank)"""
        _ck(lib().bwk_comm_barrier())

    # -- the reference's subroutine, GPU2 signature F90:525-531 (host arrays in, host arrays out)
    def biharmonic_wk_scalar(self, elem, qtens, grads, deriv, nets, nete):
        a, b = C.c_int(nets), C.c_int(nete)
        _ck(lib().bwk_biharmonic_wk_scalar(_ptr(elem), _ptr(qtens), None if grads is None else _ptr(grads), _ptr(deriv), C.byref(a), C.byref(b)))

    # -- CPU/GPU1 signature F90:186-191: qtens(np,np,nlev,qsize,nets:nete)
    def biharmonic_wk_scalar_nets(self, elem, qtens_nets, deriv, nets, nete):
        a, b = C.c_int(nets), C.c_int(nete)
        _ck(lib().bwk_biharmonic_wk_scalar_nets(_ptr(elem), _ptr(qtens_nets), _ptr(deriv), C.byref(a), C.byref(b)))

    # -- upstream HOMME full biharmonic through host arrays
    def biharmonic_wk_scalar_full(self, elem, qtens, deriv):
        _ck(lib().bwk_biharmonic_wk_scalar_full(_ptr(elem), _ptr(qtens), _ptr(deriv)))

    # -- the caller of this path in HOMME (prim_advection advance_hypervis_scalar; not in the reference, bwk_b200.h)
    def advance_hypervis_scalar(self, elem, qdp, dp, dpdiss, deriv, dt2, nu_q, hypervis_subcycle_q=1, limiter=True):
        """Qdp(np,np,nlev,qsize,nelemd) updated in place; dp, dpdiss (np,np,nlev,nelemd) Fortran order"""
        _ck(lib().bwk_advance_hypervis_scalar(_ptr(elem), _ptr(qdp), _ptr(dp), _ptr(dpdiss), _ptr(deriv), float(dt2), float(nu_q),
                                              int(hypervis_subcycle_q), 1 if limiter else 0))

    def upload_hypervis_fields(self, dp, dpdiss):
        _ck(lib().bwk_upload_hypervis_fields(_ptr(dp), _ptr(dpdiss)))

    def run_advance_hypervis_device(self, dt2, nu_q, hypervis_subcycle_q=1, limiter=True):
        _ck(lib().bwk_run_advance_hypervis(float(dt2), float(nu_q), int(hypervis_subcycle_q), 1 if limiter else 0))

    # -- upstream HOMME biharmonic_wk_dp3d with the Cartesian vlaplace_sphere_wk (not in the reference, bwk_b200.h); the
    #    KernelUtils must have been made with qsize = 4
    def biharmonic_wk_dp3d(self, elem, vec_sphere2cart, dptens, ttens, vtens, deriv, var_coef1=True, undamprrcart=True):
        """dptens, ttens (np,np,nlev,nelemd) and vtens (np,np,2,nlev,nelemd), Fortran order, hold dp3d, T and v on entry and
        the weak biharmonic on return; vec_sphere2cart (np,np,3,2,nelemd)"""
        for a in (dptens, ttens, vtens):
            if not (isinstance(a, np.ndarray) and a.flags.f_contiguous and a.dtype == np.float64):
                raise ValueError("biharmonic_wk_dp3d needs Fortran-ordered float64 arrays")
        _ck(lib().bwk_biharmonic_wk_dp3d(_ptr(elem), _ptr(vec_sphere2cart), _ptr(dptens), _ptr(ttens), _ptr(vtens), _ptr(deriv),
                                         1 if var_coef1 else 0, 1 if undamprrcart else 0))

    def upload_vec_sphere2cart(self, vec_sphere2cart):
        _ck(lib().bwk_upload_vec_sphere2cart(_ptr(vec_sphere2cart)))

    def run_biharmonic_dp3d_device(self, var_coef1=True, undamprrcart=True):
        _ck(lib().bwk_run_biharmonic_dp3d(1 if var_coef1 else 0, 1 if undamprrcart else 0))

    def set_topology_cubed_sphere(self, ne, elem_first=0):
        _ck(lib().bwk_set_topology_cubed_sphere(ne, elem_first))

    def set_topology_gdof(self, nelem_global, gdof, elem_first=0):
        g = np.ascontiguousarray(gdof, dtype=np.int64)
        _ck(lib().bwk_set_topology_gdof(nelem_global, _ptr(g), elem_first))

    def get_rspheremp(self):
        r = np.empty((NP, NP, self.nelemd), dtype=np.float64, order="F")
        _ck(lib().bwk_get_rspheremp(_ptr(r)))
        return r

    def set_rspheremp(self, r):
        r = np.asfortranarray(r, dtype=np.float64)
        _ck(lib().bwk_set_rspheremp(_ptr(r)))

    # -- save_qtens / compute_l2norm (F90:62-73) on the device copies
    def save_qtens(self):
        _ck(lib().bwk_save_qtens())

    def compute_l2norm(self):
        v = C.c_double()
        _ck(lib().bwk_compute_l2norm(C.byref(v)))
        return v.value

    def checksum(self):
        a, b = C.c_double(), C.c_double()
        _ck(lib().bwk_checksum(C.byref(a), C.byref(b)))
        return a.value, b.value

    def checksum_bits(self, elem_first_global=0):
        """partition-independent checksum of the field's bits (per-rank values add up mod 2^64)"""
        h = C.c_uint64()
        _ck(lib().bwk_checksum_bits(elem_first_global, C.byref(h)))
        return h.value

    def periodicity_defect(self, period_elems):
        v = C.c_double()
        _ck(lib().bwk_periodicity_defect(period_elems, C.byref(v)))
        return v.value

    def download_elements(self, first, count):
        out = np.empty((NP, NP, self.nlev, self.qsize, count), dtype=np.float64, order="F")
        _ck(lib().bwk_download_elements(first, count, _ptr(out)))
        return out

    def upload_elements(self, first, arr):
        a = np.asfortranarray(arr, dtype=np.float64)
        _ck(lib().bwk_upload_elements(first, a.shape[-1], _ptr(a)))

    def set_variant(self, v):
        _ck(lib().bwk_set_variant(v))

    def event_record(self, slot):
        _ck(lib().bwk_event_record(slot))

    def set_phase_event(self, slot):
        """record event `slot` between sweep 1 and the strip sweep of the following full-biharmonic calls (1 rank; -1 = off)"""
        _ck(lib().bwk_set_phase_event(slot))

    def event_elapsed_ms(self, a, b):
        v = C.c_float()
        _ck(lib().bwk_event_elapsed_ms(a, b, C.byref(v)))
        return v.value

    @staticmethod
    def launch_count():
        return lib().bwk_launch_count()

    # -- several GPUs
    @staticmethod
    def comm_unique_id():
        buf = (C.c_char * 128)()
        _ck(lib().bwk_comm_unique_id(buf))
        return bytes(buf)

    @staticmethod
    def comm_init(rank, nranks, id_bytes):
        buf = (C.c_char * 128).from_buffer_copy(id_bytes)
        _ck(lib().bwk_comm_init(rank, nranks, buf))
