"""CPU: the ARITHMETIC of the hot kernel -- codesign-kernels_b200/csrc/bwk_math.cuh, the source bwk_kernels.cu includes for the
device -- compiled for the host (tests/driver/bwk_math_host.cpp) and compared with the oracle (CPU order, F90:100-202) and with
the reference's own CPU module translated from its Fortran (oracle/ref_atm_py.py), with the tolerance the GPU tests state:
relative L2 <= 4e-16, max elementwise relative <= 1e-12 on the reference's inputs.  The kernel's folded form (one 2x2 matrix per
GLL point instead of Dinv / tensorVisc / Dinv^T / spheremp / rrearth^2 applied one after the other, explicit FMAs) is NOT
bit-identical to the CPU order -- the reference's own GPU forms are not either (1.3e-16 .. 1.4e-16, tests/golden/kat.json) --
so this file is where a change of that arithmetic shows without a GPU.  The data movement is not emulated here."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
L2_TOL, REL_TOL = 4e-16, 1e-12  # tests/test_gpu_parity.py


@pytest.fixture(scope="module")
def host():
    drv = os.path.join(ROOT, "tests", "driver")
    subprocess.run(["make", "-C", drv, "libbwk_math_host.so"], check=True, stdout=subprocess.DEVNULL)
    L = C.CDLL(os.path.join(drv, "libbwk_math_host.so"))
    L.bwk_math_host_laplace.argtypes = [C.c_void_p] * 3 + [C.c_int] * 5
    L.bwk_math_host_fold_metric.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int]

    def laplace(elem, q, dvv, nets=None, nete=None):
        nlev, qsize, n = q.shape[2:]
        assert L.bwk_math_host_laplace(elem.ctypes.data, q.ctypes.data, dvv.ctypes.data, nlev, qsize, n, nets or 1, nete or n) == 0
        return q

    laplace.lib = L
    return laplace


@pytest.mark.parametrize("nelemd,nlev,qsize", [(16, 72, 40), (1, 1, 1), (3, 5, 3), (7, 33, 1), (40, 31, 2)])
def test_kernel_arithmetic_against_the_oracle_on_the_reference_inputs(host, oracle, nelemd, nlev, qsize):
    dvv, elem, q = oracle.initialize_data(nelemd, nlev, qsize)
    want = oracle.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv)
    got = host(elem, q.copy(order="F"), dvv)
    l2, mr = oracle.compute_l2norm(got, want), oracle.max_rel(got, want)
    assert l2 <= L2_TOL and mr <= REL_TOL, (l2, mr)
    if (nelemd, nlev, qsize) == (16, 72, 40):
        # what the GPU measures on the same inputs (smoke: 1.5e-16 / 8.0e-16): same arithmetic, same size of drift, and of the
        # size of the reference's own GPU2-vs-CPU drift (1.33e-16)
        assert 5e-17 < l2 < 3e-16, l2
        # ... to the digits the GPU run printed (profiles/r3i_smoke.txt: "rel-L2 1.497e-16, max elementwise rel 8.019e-16")
        smoke = open(os.path.join(ROOT, "profiles", "r3i_smoke.txt")).read()
        assert f"rel-L2 {l2:.3e}, max elementwise rel {mr:.3e}" in smoke, (l2, mr)


def test_kernel_arithmetic_against_the_translated_reference(host, oracle):
    import ref_atm_py as ref

    if not ref.available():
        pytest.skip("oracle/_ref/libatm_f90ref.so is absent and /root/reference is not here to build it")
    dvv, elem, q = oracle.initialize_data(16, 72, 40)
    want = ref.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv)
    got = host(elem, q.copy(order="F"), dvv)
    assert oracle.compute_l2norm(got, want) <= L2_TOL and oracle.max_rel(got, want) <= REL_TOL


def test_kernel_arithmetic_on_well_conditioned_random_inputs(host, oracle):
    """Random inputs of both signs cancel heavily in the 4-term sums, so elementwise relative error is the wrong yardstick
    there; relative to the largest output of the slab's element the folded form stays within a few ulp."""
    rng = np.random.default_rng(3)
    n, nlev, qsize = 6, 9, 4
    dvv, elem, q = rng.random(16) - 0.5, rng.random((n, 144)) - 0.3, np.asfortranarray(rng.random((4, 4, nlev, qsize, n)) - 0.5)
    want = oracle.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv)
    got = host(elem, q.copy(order="F"), dvv)
    scale = np.abs(want).reshape(16, -1, order="F").max(axis=0)  # per slab
    err = np.abs(got - want).reshape(16, -1, order="F").max(axis=0)
    assert np.all(err <= 64 * np.finfo(float).eps * scale), float((err / scale).max())
    assert oracle.compute_l2norm(got, want) <= 2e-15


def test_nets_nete_and_fold_metric(host, oracle):
    dvv, elem, q = oracle.initialize_data(9, 12, 5)
    got = host(elem, q.copy(order="F"), dvv, 3, 7)
    want = oracle.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv, 3, 7)
    assert np.array_equal(got[..., :2], q[..., :2]) and np.array_equal(got[..., 7:], q[..., 7:])
    assert oracle.max_rel(got, want) <= REL_TOL
    # folded metric against its definition M = -(rrearth^2) spheremp * D * T * D^T per point, in numpy
    metric = np.empty((9, 64))
    host.lib.bwk_math_host_fold_metric(elem.ctypes.data, metric.ctypes.data, 9, 1)
    rr = oracle.rrearth()
    for e in (0, 8):
        E = elem[e]
        for pt in range(16):
            D11, D21, D12, D22 = E[pt], E[16 + pt], E[32 + pt], E[48 + pt]      # Dinv(i,j,c,a) at 16c + 32a (F90:24)
            T11, T21, T12, T22 = E[80 + pt], E[96 + pt], E[112 + pt], E[128 + pt]  # tensorVisc, same order (F90:26)
            G1x, G1y = T11 * D11 + T12 * D12, T11 * D21 + T12 * D22             # T * D^T (F90:128-133, 175-178)
            G2x, G2y = T21 * D11 + T22 * D12, T21 * D21 + T22 * D22
            f = -(rr * rr) * E[64 + pt]
            want4 = [f * (D11 * G1x + D12 * G2x), f * (D11 * G1y + D12 * G2y), f * (D21 * G1x + D22 * G2x), f * (D21 * G1y + D22 * G2y)]
            assert np.array_equal(metric[e, 4 * pt:4 * pt + 4], np.array(want4)), (e, pt)
