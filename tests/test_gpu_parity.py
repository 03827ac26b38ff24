"""GPU (-m gpu): the CUDA path through the C ABI against the CPU oracle on the same seeded
inputs.  Tolerances (stated, FP64): relative L2 <= 4e-16 and max elementwise relative <= 1e-12
on the reference's LCG inputs for the single Laplacian (the reference's own variant-vs-CPU
drift is 1.3e-16 .. 1.5e-16, atmosphere/results/atm.*); max elementwise relative <= 1e-12 for
the full biharmonic.  Nothing here reads /root/reference."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

L2_TOL = 4e-16
REL_TOL = 1e-12


def _oracle_laplace(oracle, nelemd, nlev, qsize, nets=None, nete=None):
    dvv, elem, q = oracle.initialize_data(nelemd, nlev, qsize)
    ref = oracle.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv, nets, nete)
    return q, ref


def _check(oracle, got, ref):
    l2 = oracle.compute_l2norm(got, ref)
    mr = oracle.max_rel(got, ref)
    assert l2 <= L2_TOL and mr <= REL_TOL, (l2, mr)
    return l2, mr


@pytest.mark.parametrize("variant", [0, 1])
def test_default_problem_device_resident(bwk, oracle, variant):
    """The reference program's sequence F90:552-578 on its built-in size (16 x 72 x 40)."""
    q, ref = _oracle_laplace(oracle, 16, 72, 40)
    with bwk.KernelUtils(16, 72, 40).allocate_data() as ku:
        ku.set_variant(variant)
        ku.initialize_data()
        assert np.array_equal(ku.qtens, q)
        ku.update_device()
        ku.run_device(1, 16)
        ku.update_host()
        l2, mr = _check(oracle, ku.qtens, ref)
        print(f"variant {variant}: L2 {l2:.3e} maxrel {mr:.3e}")
        assert ku.launch_count() >= 2


def test_default_problem_drop_in_call(bwk, oracle):
    """bwk_biharmonic_wk_scalar == the reference's timed region F90:572-576 (host in, host out)."""
    q, ref = _oracle_laplace(oracle, 16, 72, 40)
    with bwk.KernelUtils(16, 72, 40).allocate_data() as ku:
        ku.initialize_data()
        grads = np.zeros((4, 4, 2, 72, 40, 16), order="F")
        ku.biharmonic_wk_scalar(ku.elem, ku.qtens, grads, ku.deriv, 1, 16)
        _check(oracle, ku.qtens, ref)
        assert not grads.any()  # GPU2's scratch is accepted and left alone


def test_device_generator_is_bit_identical_to_host(bwk, oracle):
    with bwk.KernelUtils(16, 72, 40).allocate_data() as ku:
        ku.initialize_data()
        ku.initialize_data_device()
        got = ku.download_elements(0, 16)
        assert np.array_equal(got, ku.qtens)
        # and the device-generated metric gives the same operator
        _, ref = _oracle_laplace(oracle, 16, 72, 40)
        ku.run_device()
        _check(oracle, ku.download_elements(0, 16), ref)


@pytest.mark.parametrize("nelemd,nlev,qsize", [(1, 1, 1), (3, 5, 3), (2, 72, 10), (5, 128, 10), (7, 33, 1), (40, 31, 2)])
def test_ragged_sizes(bwk, oracle, nelemd, nlev, qsize):
    """nlev*qsize below / not a multiple of the 32-slab tile, single element, nlev=128."""
    q, ref = _oracle_laplace(oracle, nelemd, nlev, qsize)
    with bwk.KernelUtils(nelemd, nlev, qsize).allocate_data() as ku:
        ku.initialize_data()
        ku.biharmonic_wk_scalar(ku.elem, ku.qtens, None, ku.deriv, 1, nelemd)
        _check(oracle, ku.qtens, ref)


def test_nets_nete_subrange_and_cpu_signature(bwk, oracle):
    nelemd, nlev, qsize = 9, 12, 5
    q, ref = _oracle_laplace(oracle, nelemd, nlev, qsize, 3, 7)
    with bwk.KernelUtils(nelemd, nlev, qsize).allocate_data() as ku:
        ku.initialize_data()
        ku.biharmonic_wk_scalar(ku.elem, ku.qtens, None, ku.deriv, 3, 7)
        assert np.array_equal(ku.qtens[..., :2], q[..., :2]) and np.array_equal(ku.qtens[..., 7:], q[..., 7:])
        _check(oracle, ku.qtens, ref)
        # CPU-signature call: qtens(np,np,nlev,qsize,nets:nete) F90:188
        ku.initialize_data()
        sub = np.asfortranarray(ku.qtens[..., 2:7])
        ku.biharmonic_wk_scalar_nets(ku.elem, sub, ku.deriv, 3, 7)
        _check(oracle, sub, np.asfortranarray(ref[..., 2:7]))
        # empty range is a no-op (do ie = nets, nete with nete = nets-1)
        ku.initialize_data()
        ku.biharmonic_wk_scalar(ku.elem, ku.qtens, None, ku.deriv, 4, 3)
        assert np.array_equal(ku.qtens, q)


def test_bad_arguments_fail_loudly(bwk):
    with bwk.KernelUtils(4, 4, 4).allocate_data() as ku:
        ku.initialize_data()
        with pytest.raises(bwk.BwkError):
            ku.biharmonic_wk_scalar(ku.elem, ku.qtens, None, ku.deriv, 0, 4)
        with pytest.raises(bwk.BwkError):
            ku.biharmonic_wk_scalar(ku.elem, ku.qtens, None, ku.deriv, 1, 5)
        with pytest.raises(bwk.BwkError):
            ku.run_biharmonic_device()  # no topology


def test_pinned_host_arrays_and_chunked_pipeline(bwk, oracle):
    """ne=8 sized problem through pinned host memory: the e2e path splits into several chunks."""
    nelemd, nlev, qsize = 384, 72, 10
    q, ref = _oracle_laplace(oracle, nelemd, nlev, qsize)
    with bwk.KernelUtils(nelemd, nlev, qsize).allocate_data(pinned=True) as ku:
        ku.initialize_data()
        ku.biharmonic_wk_scalar(ku.elem, ku.qtens, None, ku.deriv, 1, nelemd)
        _check(oracle, ku.qtens, ref)


def test_linearity_and_idempotent_reinit(bwk, oracle):
    """Size-independent properties: L(a*x) == a*L(x) for a power of two (exact), and two runs
    from re-initialised data are bit-identical (deterministic kernel, no atomics)."""
    with bwk.KernelUtils(8, 16, 4).allocate_data() as ku:
        ku.initialize_data()
        ku.update_device()
        ku.run_device()
        ku.update_host()
        a = ku.qtens.copy(order="F")
        ku.initialize_data()
        ku.qtens *= 4.0
        ku.update_device()
        ku.run_device()
        ku.update_host()
        assert np.array_equal(ku.qtens, 4.0 * a)
        ku.initialize_data()
        ku.update_device()
        ku.run_device()
        ku.update_host()
        assert np.array_equal(ku.qtens, a)


def test_checksum_bits_is_position_aware_and_partition_independent(bwk):
    """bwk_checksum_bits == the same formula in numpy; two half-size contexts holding the two halves of the field add up
    (mod 2^64) to the whole; swapping two values changes it."""
    def ref(q, index0):
        x = q.ravel(order="F").view(np.uint64).copy()
        with np.errstate(over="ignore"):
            x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
            x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
            x = x ^ (x >> np.uint64(31))
            w = np.uint64(2) * (np.uint64(index0) + np.arange(x.size, dtype=np.uint64)) + np.uint64(1)
            return int(np.sum(x * w, dtype=np.uint64))
    n, nlev, qsize = 10, 7, 3
    with bwk.KernelUtils(n, nlev, qsize).allocate_data() as ku:
        ku.initialize_data()
        ku.update_device()
        whole = ku.checksum_bits(0)
        assert whole == ref(ku.qtens, 0)
        q = ku.qtens.copy(order="F")
        ku.qtens[0, 0, 0, 0, 0], ku.qtens[1, 0, 0, 0, 0] = q[1, 0, 0, 0, 0], q[0, 0, 0, 0, 0]
        ku.update_device(elem=False, deriv=False)
        assert ku.checksum_bits(0) != whole
    parts = []
    for first, cnt in ((0, 4), (4, 6)):
        with bwk.KernelUtils(cnt, nlev, qsize).allocate_data() as ku:
            ku.qtens[...] = q[..., first:first + cnt]
            ku.update_device(elem=False, deriv=False)
            parts.append(ku.checksum_bits(first))
    assert (parts[0] + parts[1]) % (1 << 64) == whole


def test_high_entropy_inputs(bwk, oracle):
    """Robustness run: splitmix64 values instead of the 17-bit LCG lattice.
    All operands stay positive, so the same tolerances hold."""
    nelemd, nlev, qsize = 32, 24, 3
    with bwk.KernelUtils(nelemd, nlev, qsize).allocate_data() as ku:
        ku.initialize_data_device(seed_mode=1)
        q = ku.download_elements(0, nelemd)
        # regenerate the same values on the host: splitmix64(d) / 2^53
        def sm(d):
            x = (d.astype(np.uint64) + np.uint64(0x9E3779B97F4A7C15))
            x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
            x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
            x = x ^ (x >> np.uint64(31))
            return (x >> np.uint64(11)).astype(np.float64) * 2.0 ** -53
        dvv = sm(np.arange(16))
        elem = sm(16 + np.arange(144 * nelemd)).reshape(nelemd, 144)
        qh = sm(16 + 144 * nelemd + np.arange(q.size)).reshape(q.shape, order="F")
        assert np.array_equal(qh, q)
        ref = oracle.biharmonic_wk_scalar_cpu(elem, qh.copy(order="F"), dvv)
        ku.run_device()
        got = ku.download_elements(0, nelemd)
        assert oracle.max_rel(got, ref) <= REL_TOL and oracle.compute_l2norm(got, ref) <= L2_TOL


def test_ne30_q10_subset_against_oracle(bwk, oracle):
    """BASELINE config 2 (ne30np4, nlev 72, qsize 10) at full size on the device; the oracle
    checks the first and last 24 elements, a checksum pins the rest to the periodic structure."""
    n, nlev, qsize = 5400, 72, 10
    with bwk.KernelUtils(n, nlev, qsize).allocate_data(host_arrays=False) as ku:
        ku.initialize_data_device()
        ku.run_device()
        for first in (0, n - 24, 2700):
            dvv, elem, q = oracle.initialize_subset(n, nlev, qsize, first, 24)
            ref = oracle.biharmonic_wk_scalar_cpu(elem, q, dvv)
            _check(oracle, ku.download_elements(first, 24), ref)


def test_ne120_q40_full_size_properties(bwk, oracle):
    """BASELINE's headline size (ne120np4 L72 q40, 86,400 elements, 31.85 GB in place).
    The LCG inputs repeat every 8192 elements (tests/test_oracle_kat.py), so the output must be
    exactly periodic too (bitwise), and a window straddling the first period boundary must match
    the oracle."""
    n, nlev, qsize = 86400, 72, 40
    with bwk.KernelUtils(n, nlev, qsize).allocate_data(host_arrays=False) as ku:
        ku.initialize_data_device()
        assert ku.periodicity_defect(8192) == 0.0
        ku.run_device()
        ku.sync()
        assert ku.periodicity_defect(8192) == 0.0
        for first in (0, 8190, n - 4):
            dvv, elem, q = oracle.initialize_subset(n, nlev, qsize, first, 4)
            ref = oracle.biharmonic_wk_scalar_cpu(elem, q, dvv)
            _check(oracle, ku.download_elements(first, 4), ref)
        s, ss = ku.checksum()
        assert s < 0 and ss > 0 and np.isfinite(s) and np.isfinite(ss)


def test_ne30_q40_whole_array_against_the_oracle(bwk, oracle):
    """BASELINE config 3 verbatim: ne30np4 nlev=72 qsize=40 (5,400 elements, 1.99 GB), "GPU result compared against the
    reference CPU output" on the WHOLE array -- the single Laplacian the reference computes (F90:186-200) and the full
    biharmonic.  The CPU side is the oracle's OpenMP build (bit-identical to the parity build,
    tests/test_oracle_kat.py::test_omp_*)."""
    ne, nlev, qsize = 30, 72, 40
    n = 6 * ne * ne
    dvv, elem, q = oracle.initialize_data(n, nlev, qsize)
    with bwk.KernelUtils(n, nlev, qsize).allocate_data(host_arrays=False) as ku:
        ku.initialize_data_device()
        ku.set_topology_cubed_sphere(ne)
        assert np.array_equal(ku.download_elements(0, n), q)  # device generator == sequential host generator, whole array
        ku.run_device()
        got = ku.download_elements(0, n)
        ref = oracle.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv, omp=True)
        l2, mr = _check(oracle, got, ref)
        print(f"ne30 q40 single Laplacian, whole array: L2 {l2:.3e} maxrel {mr:.3e}")
        ku.initialize_data_device()
        ku.run_biharmonic_device()
        got = ku.download_elements(0, n)
        dss = oracle.Dss(n, oracle.cubed_sphere_gdof(ne))
        ref = dss.biharmonic_full(elem, dss.rspheremp(elem), q, dvv, omp=True)
        mr = oracle.max_rel(got, ref)
        print(f"ne30 q40 full biharmonic, whole array: maxrel {mr:.3e}")
        assert mr <= REL_TOL, mr


def test_ne120_q40_full_biharmonic_headline_config(bwk, oracle):
    """The configuration bench.py times (BASELINE config 4 on one GPU: ne120np4 L72 q40, full biharmonic, 63.7 GB with
    the sweep-1 buffer).  Elements at cube corners, on face edges, next to the 8,192-element period boundary of the
    LCG inputs and in face interiors are compared with the patch oracle (bit-identical to the whole-sphere oracle,
    tests/test_plan.py::test_patch_oracle_equals_full_oracle); face-interior elements 8,192 apart see identical inputs
    in identical positions and must agree bit for bit; two runs give the same checksum."""
    from patch_oracle import patch_oracle_full

    ne, nlev, qsize = 120, 72, 40
    n, f = 6 * ne * ne, ne * ne
    gd = oracle.cubed_sphere_gdof(ne).reshape(n, 16)
    with bwk.KernelUtils(n, nlev, qsize).allocate_data(host_arrays=False) as ku:
        ku.initialize_data_device()
        ku.set_topology_cubed_sphere(ne)
        ku.run_biharmonic_device()
        ku.sync()
        picks = [0, ne - 1, f - 1, n - 1,  # cube corners (3 elements share the vertex)
                 60, 60 * ne, 2 * f + 119 * ne + 17, 5 * f + 33 * ne + ne - 1,  # face edges
                 8191, 8192, 3 * f + 60 * ne + 60, 4 * f + 7 * ne + 101]  # period boundary, interiors
        worst = 0.0
        for e in picks:
            ref = patch_oracle_full(oracle, ne, nlev, qsize, gd, e)
            mr = oracle.max_rel(ku.download_elements(e, 1), ref)
            worst = max(worst, mr)
            assert mr <= REL_TOL, (e, mr)
        print(f"ne120 q40 full biharmonic vs patch oracle at {len(picks)} elements: worst maxrel {worst:.3e}")
        for e in (5 * ne + 7, 2 * f + 9 * ne + 3, 5 * f + 40 * ne + 50):
            assert np.array_equal(ku.download_elements(e, 1), ku.download_elements(e + 8192, 1)), e
        c1 = ku.checksum()
        ku.initialize_data_device()
        ku.run_biharmonic_device()
        ku.sync()
        assert ku.checksum() == c1 and np.isfinite(c1[0]) and c1[1] > 0


# ------------------------------------------------------------------ full biharmonic (parity unpinned)
def _oracle_full(oracle, ne, nlev, qsize):
    n = 6 * ne * ne
    dvv, elem, q = oracle.initialize_data(n, nlev, qsize)
    dss = oracle.Dss(n, oracle.cubed_sphere_gdof(ne))
    rs = dss.rspheremp(elem)
    ref = dss.biharmonic_full(elem, rs, q.copy(order="F"), dvv)
    return q, rs, ref


@pytest.mark.parametrize("ne,nlev,qsize", [(1, 4, 2), (2, 8, 3), (3, 72, 2), (4, 33, 1), (8, 16, 4)])
def test_full_biharmonic_small_cubed_spheres(bwk, oracle, ne, nlev, qsize):
    n = 6 * ne * ne
    q, rs, ref = _oracle_full(oracle, ne, nlev, qsize)
    with bwk.KernelUtils(n, nlev, qsize).allocate_data() as ku:
        ku.initialize_data()
        ku.set_topology_cubed_sphere(ne)
        ku.update_device()
        assert np.array_equal(ku.get_rspheremp(), rs)
        ku.biharmonic_wk_scalar_full(ku.elem, ku.qtens, ku.deriv)
        mr = oracle.max_rel(ku.qtens, ref)
        assert mr <= REL_TOL, mr
        # device-resident entry gives the same bits as the host-buffer entry
        a = ku.qtens.copy(order="F")
        ku.initialize_data()
        ku.update_device()
        ku.run_biharmonic_device()
        ku.update_host()
        assert np.array_equal(ku.qtens, a)


def test_full_biharmonic_ne30_one_tracer(bwk, oracle):
    """ne30 topology (5,400 elements, 48,602 unique nodes) with a single tracer so that the CPU
    oracle finishes in seconds; qsize only multiplies independent slabs."""
    ne, nlev, qsize = 30, 72, 1
    n = 6 * ne * ne
    q, rs, ref = _oracle_full(oracle, ne, nlev, qsize)
    with bwk.KernelUtils(n, nlev, qsize).allocate_data(host_arrays=False) as ku:
        ku.initialize_data_device()
        ku.set_topology_cubed_sphere(ne)
        ku.run_biharmonic_device()
        got = ku.download_elements(0, n)
        mr = oracle.max_rel(got, ref)
        assert mr <= REL_TOL, mr


def test_full_biharmonic_constant_field_invariants(bwk, oracle):
    """DSS invariants without the oracle: with rspheremp from the same DSS, a field that is
    already continuous (constant) is left unchanged by sum*rspheremp only if spheremp were the
    weights; here we check the cheaper exact one: zero in -> zero out, and scaling by 2 is exact."""
    ne, nlev, qsize = 4, 8, 2
    n = 6 * ne * ne
    with bwk.KernelUtils(n, nlev, qsize).allocate_data() as ku:
        ku.initialize_data()
        ku.set_topology_cubed_sphere(ne)
        ku.update_device()
        ku.run_biharmonic_device()
        ku.update_host()
        a = ku.qtens.copy(order="F")
        ku.initialize_data()
        ku.qtens *= 2.0
        ku.update_device(elem=False, deriv=False)
        ku.run_biharmonic_device()
        ku.update_host()
        assert np.array_equal(ku.qtens, 2.0 * a)
        ku.qtens[...] = 0.0
        ku.update_device(elem=False, deriv=False)
        ku.run_biharmonic_device()
        ku.update_host()
        assert not ku.qtens.any()
