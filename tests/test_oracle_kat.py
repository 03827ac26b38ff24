"""CPU: pins the C oracle against the reference's own published known-answer scalars
(tests/golden/kat.json, transcribed from atmosphere/results/* by tests/golden/make_kat.py)."""
import json
import os

import numpy as np

KAT = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "kat.json")))


def _default(oracle):
    c = KAT["config"]
    return oracle.initialize_data(c["nelemd"], c["nlev"], c["qsize"])


def test_rrearth_is_float32_rounded(oracle):
    # F90:14 has no kind suffix -> default real
    assert oracle.rrearth() == float(np.float32(0.00000016666666666666))
    assert oracle.rrearth() == KAT["survey_derived"]["rrearth"]
    assert oracle.rrearth() == float.fromhex("0x1.65e9f8p-23")


def test_lcg_inputs(oracle):
    dvv, elem, q = _default(oracle)
    sd = KAT["survey_derived"]
    ints = [int(round(v * 131072)) for v in list(dvv[:6])]
    assert ints == sd["lcg_first_ints"]
    assert list(dvv[:3]) == sd["dvv_1_123"]
    assert np.allclose(q[:3, 0, 0, 0, 0], sd["qtens_first3"], atol=5e-9)
    assert q[3, 3, 71, 39, 15] == sd["qtens_last"]
    # Dinv(1,1,1,1) of element 1 is draw 17
    assert elem[0, 0] == ((1301 * int(dvv[15] * 131072) + 97) % 131072) / 131072.0


def test_kat1_gpu2_order_vs_cpu_order(oracle):
    dvv, elem, q = _default(oracle)
    ref = oracle.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv)
    g2 = oracle.biharmonic_wk_scalar_gpu2(elem, q.copy(order="F"), dvv)
    l2 = oracle.compute_l2norm(g2, ref)
    assert abs(l2 - KAT["KAT1_gpu2_vs_cpu_rel_l2"]) <= 1e-13 * l2, l2  # observed: all 17 digits


def test_kat2_gpu1_order_vs_cpu_order(oracle):
    dvv, elem, q = _default(oracle)
    ref = oracle.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv)
    g1 = oracle.biharmonic_wk_scalar_gpu1(elem, q.copy(order="F"), dvv)
    l2 = oracle.compute_l2norm(g1, ref)
    assert abs(l2 - KAT["KAT2_gpu1_vs_cpu_rel_l2"]) <= 1e-13 * l2, l2


def test_survey_cross_check_values(oracle):
    dvv, elem, q = _default(oracle)
    out = oracle.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv)
    sd = KAT["survey_derived"]
    assert np.allclose(out[:, 0, 0, 0, 0], sd["out_i_1_1_1_1"], rtol=1e-12, atol=0)
    assert np.allclose(out[0, :, 0, 0, 0], sd["out_1_j_1_1_1"], rtol=1e-12, atol=0)
    assert np.isclose(out[3, 3, 71, 39, 15], sd["out_last"], rtol=1e-12, atol=0)
    assert np.isclose(out.min(), sd["min_out"], rtol=1e-12) and np.isclose(out.max(), sd["max_out"], rtol=1e-12)
    import math

    assert np.isclose(math.fsum(out.ravel()), sd["sum_out"], rtol=1e-12)
    assert np.isclose(math.fsum((out.ravel() ** 2)), sd["sum_out_sq"], rtol=1e-12)


def test_nets_nete_subrange(oracle):
    dvv, elem, q = oracle.initialize_data(6, 8, 3)
    full = oracle.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv)
    part = oracle.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv, nets=2, nete=4)
    assert np.array_equal(part[..., 1:4], full[..., 1:4])
    assert np.array_equal(part[..., 0], q[..., 0]) and np.array_equal(part[..., 4:], q[..., 4:])


def test_subset_generator_matches_sequential(oracle):
    dvv, elem, q = oracle.initialize_data(40, 5, 3)
    dvv2, elem2, q2 = oracle.initialize_subset(40, 5, 3, 7, 9)
    assert np.array_equal(dvv, dvv2)
    assert np.array_equal(elem[7:16], elem2)
    assert np.array_equal(q[..., 7:16], q2)


def test_omp_variant_is_bit_identical(oracle):
    dvv, elem, q = oracle.initialize_data(12, 9, 4)
    a = oracle.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv)
    b = oracle.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv, omp=True)
    assert np.array_equal(a, b)


def test_lcg_period_makes_elements_repeat_every_8192(oracle):
    # the size-independent property the full-size GPU test relies on
    n = 86400
    _, e0, q0 = oracle.initialize_subset(n, 3, 2, 5, 2)
    _, e1, q1 = oracle.initialize_subset(n, 3, 2, 5 + 8192, 2)
    assert np.array_equal(e0, e1)
    # qtens offset 8192 elements * 16*nlev*qsize doubles is a multiple of the LCG period only when
    # 16*nlev*qsize*8192 % 131072 == 0, i.e. always (8192*16 = 131072)
    assert np.array_equal(q0, q1)


def test_omp_full_biharmonic_is_bit_identical(oracle):
    """the timing build (-O3 -fopenmp) of Laplace -> DSS -> rspheremp -> Laplace gives the parity build's bits, so the
    whole-array ne30 q40 GPU test (BASELINE config 3) may use it"""
    ne, nlev, qsize = 3, 5, 2
    n = 6 * ne * ne
    dvv, elem, q = oracle.initialize_data(n, nlev, qsize)
    dss = oracle.Dss(n, oracle.cubed_sphere_gdof(ne))
    rs = dss.rspheremp(elem)
    a = dss.biharmonic_full(elem, rs, q.copy(order="F"), dvv)
    b = dss.biharmonic_full(elem, rs, q.copy(order="F"), dvv, omp=True)
    assert np.array_equal(a, b)


def test_parallel_closed_form_initialize_data_equals_the_sequential_one(oracle):
    dvv, elem, q = oracle.initialize_data(37, 5, 3)
    dvv2, elem2, q2 = oracle.initialize_data_parallel(37, 5, 3, threads=3)
    assert np.array_equal(dvv, dvv2) and np.array_equal(elem, elem2) and np.array_equal(q, q2)
