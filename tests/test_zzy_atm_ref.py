"""oracle/_ref for the hot path itself: the reference's three implementations of biharmonic_wk_scalar
(atmosphere/biharmonic_wk_kernel.F90, cited F90:<line>) translated mechanically from the Fortran by oracle/f90_to_c.py (unit
mode) and compiled here -- see oracle/ref_atm_py.py; this image has no Fortran compiler.

  * the translations reproduce the reference's PUBLISHED outputs: `GPU2 L2 norm` and `GPU1 L2 norm` of atmosphere/results/atm.*
    (tests/golden/kat.json) to all 17 digits -- numbers that are pure rounding noise and so depend on every operation's order
  * oracle/bwk_oracle.c (CPU, GPU1 and GPU2 orders) == the translations BIT FOR BIT, on the LCG inputs, random inputs, ragged
    sizes and nets:nete sub-ranges
  * (-m gpu) the CUDA path against the translated CPU form directly, with the tolerance of tests/test_gpu_parity.py

This file sorts late on purpose: its GPU test is new.  Nothing here reads /root/reference at run time when oracle/_ref travelled.
"""
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
KAT = json.load(open(os.path.join(ROOT, "tests", "golden", "kat.json")))


@pytest.fixture(scope="module")
def ref():
    import ref_atm_py

    if not ref_atm_py.available():
        pytest.skip("oracle/_ref/libatm_f90ref.so is absent and /root/reference is not here to build it")
    try:
        ref_atm_py.lib()
    except OSError as e:  # the checker itself could not be loaded on this box: nothing was compared
        pytest.skip(f"oracle/_ref library did not load: {e}")
    return ref_atm_py


def _bits(a):
    return a.view(np.uint64)


def _random_inputs(nelemd, nlev, qsize, seed):
    rng = np.random.default_rng(seed)
    return rng.random(16) - 0.5, rng.random((nelemd, 144)) - 0.3, np.asfortranarray(rng.random((4, 4, nlev, qsize, nelemd)) - 0.5)


def test_translated_reference_reproduces_the_published_norms(oracle, ref):
    """The reference program's two known-answer lines (F90:569 commented GPU1 block, F90:578) from the translated modules on the
    reference's own inputs; the inputs and the norm (F90:48-58, 69-73, 77-91) are the oracle's restatements, themselves pinned by
    tests/test_oracle_kat.py.  All 17 digits, as printed by gfortran 4.9.3, gfortran 7.1.0 and pgfortran 17.7."""
    c = KAT["config"]
    dvv, elem, q = oracle.initialize_data(c["nelemd"], c["nlev"], c["qsize"])
    cpu = ref.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv)
    g1 = ref.biharmonic_wk_scalar_gpu1(elem, q.copy(order="F"), dvv)
    g2 = ref.biharmonic_wk_scalar_gpu2(elem, q.copy(order="F"), dvv)
    kat1, kat2 = oracle.compute_l2norm(g2, cpu), oracle.compute_l2norm(g1, cpu)
    assert "%.16E" % kat1 == "%.16E" % KAT["KAT1_gpu2_vs_cpu_rel_l2"] == "1.3341964743086566E-16"
    assert "%.16E" % kat2 == "%.16E" % KAT["KAT2_gpu1_vs_cpu_rel_l2"] == "1.4477731365593603E-16"
    # a double-precision rrearth (or one reordered sum) moves these in the second digit: the check is sharp
    assert not np.array_equal(_bits(cpu), _bits(g1)) and not np.array_equal(_bits(cpu), _bits(g2))


@pytest.mark.parametrize("nelemd,nlev,qsize", [(16, 72, 40), (1, 1, 1), (3, 5, 3), (2, 72, 10), (5, 128, 10), (7, 33, 1), (40, 31, 2)])
def test_oracle_cpu_order_is_bit_identical_to_the_translated_cpu_module(oracle, ref, nelemd, nlev, qsize):
    """oracle/bwk_oracle.c bwko_biharmonic_wk_scalar_cpu against module biharmonic_wk_scalar_cpu (F90:100-202): the reference's
    inputs at its own size (F90:11-17) and at the ragged sizes of tests/test_gpu_parity.py, then random inputs of both signs."""
    for dvv, elem, q in (oracle.initialize_data(nelemd, nlev, qsize), _random_inputs(nelemd, nlev, qsize, 11 * nelemd + nlev)):
        a = oracle.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv)
        b = ref.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv)
        assert np.isfinite(b).all() and not np.array_equal(b, q)
        assert np.array_equal(_bits(a), _bits(b))


def test_nets_nete_subrange_matches_and_leaves_the_rest_alone(oracle, ref):
    dvv, elem, q = _random_inputs(9, 12, 5, 3)
    a = oracle.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv, 3, 7)
    b = ref.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv, 3, 7)
    assert np.array_equal(_bits(a), _bits(b))
    assert np.array_equal(b[..., :2], q[..., :2]) and np.array_equal(b[..., 7:], q[..., 7:]) and not np.array_equal(b[..., 2:7], q[..., 2:7])


@pytest.mark.parametrize("nelemd,nlev,qsize", [(16, 72, 40), (3, 16, 2), (2, 20, 3)])
def test_oracle_gpu_orders_are_bit_identical_to_the_translated_gpu_modules(oracle, ref, nelemd, nlev, qsize):
    """The GPU1 (F90:209-362) and GPU2 (F90:369-536) operation orders of the oracle exist only to reproduce the known-answer
    norms; they equal the translated modules bit for bit too.  nlev = 20: GPU1 walks nlev/kchunk = 2 whole blocks of 8 levels and
    leaves levels 17..20 untouched (F90:331), GPU2 clamps its last block (F90:409, 458)."""
    for dvv, elem, q in (oracle.initialize_data(nelemd, nlev, qsize), _random_inputs(nelemd, nlev, qsize, 5)):
        a1 = oracle.biharmonic_wk_scalar_gpu1(elem, q.copy(order="F"), dvv)
        b1 = ref.biharmonic_wk_scalar_gpu1(elem, q.copy(order="F"), dvv)
        assert np.array_equal(_bits(a1), _bits(b1))
        if nlev % 8:
            assert np.array_equal(b1[:, :, nlev // 8 * 8:], q[:, :, nlev // 8 * 8:])
        a2 = oracle.biharmonic_wk_scalar_gpu2(elem, q.copy(order="F"), dvv)
        b2 = ref.biharmonic_wk_scalar_gpu2(elem, q.copy(order="F"), dvv)
        assert np.isfinite(b2).all() and np.array_equal(_bits(a2), _bits(b2))


@pytest.mark.gpu
@pytest.mark.parametrize("nelemd,nlev,qsize", [(16, 72, 40), (7, 33, 1)])
def test_gpu_against_the_translated_reference(bwk, oracle, ref, nelemd, nlev, qsize):
    """The CUDA path through the drop-in call (F90:525-534 signature, host arrays) against the translated CPU module, no
    hand-written oracle in between.  Tolerance as tests/test_gpu_parity.py: relative L2 <= 4e-16, max elementwise relative
    <= 1e-12 (the folded-metric kernel is not bit-identical to the CPU order; the reference's own GPU forms differ by 1.3e-16)."""
    with bwk.KernelUtils(nelemd, nlev, qsize).allocate_data() as ku:
        ku.initialize_data()
        dvv = np.ascontiguousarray(ku.deriv["Dvv"]).reshape(16)
        elem = np.ascontiguousarray(ku.elem.view(np.float64).reshape(nelemd, 144))
        want = ref.biharmonic_wk_scalar_cpu(elem, ku.qtens.copy(order="F"), dvv)
        ku.biharmonic_wk_scalar(ku.elem, ku.qtens, None, ku.deriv, 1, nelemd)
        l2, mr = oracle.compute_l2norm(ku.qtens, want), oracle.max_rel(ku.qtens, want)
        assert l2 <= 4e-16 and mr <= 1e-12, (l2, mr)


# ------------------------------------------------------------------------------------------------ the translator's unit mode
UNIT_DEMO = """
! synthetic modules for the translator test (not reference code)
module consts
  implicit none
  integer, parameter :: rk = selected_real_kind(13)
  integer, parameter :: n = 3
  integer, parameter :: m = 4
  real(kind=rk), parameter :: third = 0.3333333
  integer :: count = 5
  type :: pack_t
    real(kind=rk) :: w(n,2)
    real(kind=rk) :: bias
  end type pack_t
end module consts

module work
  use consts
  implicit none
  private
  public :: driver
contains
  function scaled(x, p) result(y)
    real(kind=rk), intent(in) :: x(n)
    type (pack_t), intent(in) :: p
    real(kind=rk)             :: y(n)
    integer i
    do i = 1 , n
      y(i) = x(n+1-i) * p%w(i,2) + p%bias * third
    enddo
  end function scaled

  subroutine accumulate(a, packs, total, k)
    real(kind=rk), intent(in   ) :: a(n,m)
    type (pack_t), intent(in   ) :: packs(:)
    real(kind=rk), intent(inout) :: total(m)
    integer      , intent(in   ) :: k
    real(kind=rk) :: col(n)
    col = a(:,k)
    total(k) = total(k) + sum(col(:)*packs(k)%w(:,1))
  end subroutine accumulate

  subroutine driver(a, packs, total)
    real(kind=rk), intent(inout) :: a(n,m)
    type (pack_t), intent(in   ) :: packs(:)
    real(kind=rk), intent(  out) :: total(m)
    real(kind=rk) :: keep(n,m)
    integer :: k
    keep = a
    do k = 1 , m
      total(k) = 0
      a(:,k) = scaled(a(:,k), packs(k))
      call accumulate(keep, packs(:), total, k)
      if (k > count) total(k) = -1.
    enddo
  end subroutine driver
end module work
"""


def test_translator_unit_mode_semantics_on_synthetic_modules(tmp_path):
    import ctypes as C
    import subprocess

    src, out, so = tmp_path / "u.F90", tmp_path / "u.c", tmp_path / "libu.so"
    src.write_text(UNIT_DEMO)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "f90_to_c.py"), str(src), "driver", "--module", "work", "--uses", "consts",
                        "--runtime", "count", "--export", "demo_driver", "--out", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "0.3333333f" in out.read_text()  # a real parameter keeps the precision of its literal
    subprocess.run(["/usr/bin/gcc", "-O2", "-ffp-contract=off", "-fPIC", "-std=gnu11", "-shared", "-o", str(so), str(out), "-lm"], check=True)
    L = C.CDLL(str(so))
    n, m = 3, 4
    rng = np.random.default_rng(1)
    a = np.asfortranarray(rng.random((n, m)) - 0.5)
    packs = rng.random((m, 2 * n + 1)) - 0.5  # pack_t = w(n,2) column-major, then bias
    total = np.full(m, np.nan)
    a0 = a.copy(order="F")
    L.demo_driver.argtypes = [C.c_int] + [C.c_void_p] * 3
    assert L.demo_driver(3, a.ctypes.data, packs.ctypes.data, total.ctypes.data) == 0  # count = 3: total(4) = -1
    third = np.float64(np.float32(0.3333333))
    for k in range(m):
        w, bias = packs[k, :2 * n].reshape((n, 2), order="F"), packs[k, 2 * n]
        want = np.array([a0[n - 1 - i, k] * w[i, 1] + bias * third for i in range(n)])  # reads the whole old column first
        assert np.array_equal(a[:, k].view(np.uint64), want.view(np.uint64))
        s = 0.0
        for i in range(n):
            s = s + a0[i, k] * w[i, 0]
        assert total[k] == (0.0 + s if k < 3 else -1.0)
    # outside the subset -> refused
    bad = UNIT_DEMO.replace("col = a(:,k)", "col = matmul(a, total)")
    src.write_text(bad)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "f90_to_c.py"), str(src), "driver", "--module", "work", "--uses", "consts",
                        "--out", str(out)], capture_output=True, text=True)
    assert r.returncode != 0 and "f90_to_c:" in r.stderr


def test_baseline_config_3_whole_array_oracle_equals_the_translated_reference(oracle, ref):
    """BASELINE.json config 3 verbatim -- ne30np4 (5,400 elements) nlev = 72 qsize = 40, 15.5 M slabs, 2 GB -- and config 2
    (qsize = 10): the oracle (OpenMP build, the one bench.py and the large GPU tests use) against the translated CPU module on
    the WHOLE array, bit for bit.  The GPU test test_ne30_q40_whole_array_against_the_oracle compares the CUDA result with this
    oracle at the same size."""
    for qsize in (10, 40):
        dvv, elem, q = oracle.initialize_data_parallel(5400, 72, qsize)
        a = oracle.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv, omp=True)
        b = ref.biharmonic_wk_scalar_cpu(elem, q, dvv)  # in place on q: 2 GB are enough once
        assert np.array_equal(_bits(a), _bits(b))
        del a, b, q


def test_oracle_hashes_to_the_committed_digests_of_the_translated_reference(oracle):
    """tests/golden/atm_translated.json (made by tests/golden/make_atm_golden.py from the translated reference) against the
    hand-written oracle: runs anywhere, with or without oracle/_ref.  The default case also carries the two published norms."""
    import hashlib

    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "atm_translated.json")))
    sha = lambda a: hashlib.sha256(np.asfortranarray(a).tobytes(order="F")).hexdigest()  # noqa: E731
    for c in gold["cases"]:
        dvv, elem, q = oracle.initialize_data(c["nelemd"], c["nlev"], c["qsize"])
        assert sha(q) == c["input_sha256"]
        cpu = oracle.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv)
        g1 = oracle.biharmonic_wk_scalar_gpu1(elem, q.copy(order="F"), dvv)
        g2 = oracle.biharmonic_wk_scalar_gpu2(elem, q.copy(order="F"), dvv)
        assert (sha(cpu), sha(g1), sha(g2)) == (c["cpu_sha256"], c["gpu1_sha256"], c["gpu2_sha256"])
    d = gold["cases"][0]
    assert (d["nelemd"], d["nlev"], d["qsize"]) == (16, 72, 40)
    assert d["gpu2_vs_cpu_rel_l2"] == "%.16E" % KAT["KAT1_gpu2_vs_cpu_rel_l2"] and d["gpu1_vs_cpu_rel_l2"] == "%.16E" % KAT["KAT2_gpu1_vs_cpu_rel_l2"]
    n = gold["nets_nete_case"]
    dvv, elem, q = oracle.initialize_data(n["nelemd"], n["nlev"], n["qsize"])
    assert sha(oracle.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv, n["nets"], n["nete"])) == n["cpu_sha256"]


STATE_DEMO_VARS = """
! synthetic module for the translator's state mode (not reference code)
#define UNUSED_MACRO(x) x
module shelf
  implicit none
  public
  integer, parameter :: RK = selected_real_kind(12)
  integer :: n, m, &! sizes, set by the caller
     j, i        ! loop indices are module variables too
  namelist /sizes/ n, m
  real (kind=RK), parameter :: gain = 0.1
#ifdef WIDE
  real (kind=RK), dimension(:,:), allocatable :: acc
#else
  real (kind=RK), dimension(:), allocatable :: acc
#endif
  real (kind=RK), dimension(:,:), allocatable :: &
     field, &! (n, m)
     weight
  integer, dimension(:), allocatable :: first
contains
  subroutine alloc_all
    allocate(field(n,m), weight(n,m), &
             first(m))
#ifdef WIDE
    allocate(acc(n,m))
#else
    allocate(acc(n))
#endif
  end subroutine alloc_all
end module shelf
"""
STATE_DEMO_MAIN = """
program main
  use shelf
  implicit none
  call alloc_all
  call fold
contains
  subroutine fold
    do j = 1, m
      do i = 1, n
        acc(i) = 0.0_RK
      end do
      do i = first(j), n
        acc(i) = acc(i) + field(i,j)*weight(i,j) * gain + sign(1.0_RK, field(i,j))
      end do
      do i = 1, n
        field(i,j) = acc(i)
      end do
    end do
  end subroutine fold
end program main
"""


def test_translator_state_mode_on_synthetic_sources(tmp_path):
    """Module state as file-scope variables, #ifdef with nothing defined, shapes from the module's allocate statements, several
    source files, routines without an argument list, dimension(..) attributes, sign -- the constructs of nested_loops."""
    import ctypes as C
    import subprocess

    (tmp_path / "shelf.F90").write_text(STATE_DEMO_VARS)
    (tmp_path / "main.F90").write_text(STATE_DEMO_MAIN)
    out, so = tmp_path / "s.c", tmp_path / "libs.so"
    r = subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "f90_to_c.py"), str(tmp_path / "main.F90"), "fold", "--also", str(tmp_path / "shelf.F90"),
                        "--module", "main", "--uses", "shelf", "--state", "--export", "demo", "--out", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    src = out.read_text()
    assert "0.1f" in src and "static double *acc;" in src and "acc_n1" not in src  # the #else branch: acc is rank 1
    subprocess.run(["/usr/bin/gcc", "-O2", "-ffp-contract=off", "-fPIC", "-std=gnu11", "-shared", "-o", str(so), str(out), "-lm"], check=True)
    L = C.CDLL(str(so))
    L.demo_set_int.argtypes, L.demo_bind.argtypes = [C.c_char_p, C.c_int], [C.c_char_p, C.c_void_p]
    n, m = 5, 3
    rng = np.random.default_rng(2)
    field, weight = np.asfortranarray(rng.random((n, m)) - 0.5), np.asfortranarray(rng.random((n, m)))
    first, acc = np.array([1, 3, 6], dtype=np.int32), np.full(n, np.nan)
    f0 = field.copy(order="F")
    assert L.demo_set_int(b"n", n) == 0 and L.demo_set_int(b"m", m) == 0 and L.demo_set_int(b"nope", 1) == 1
    for name, a in (("field", field), ("weight", weight), ("first", first), ("acc", acc)):
        assert L.demo_bind(name.encode(), a.ctypes.data) == 0
    assert L.demo_fold() == 0
    gain = np.float64(np.float32(0.1))
    for j in range(m):
        want = np.zeros(n)
        for i in range(first[j] - 1, n):
            want[i] = 0.0 + f0[i, j] * weight[i, j] * gain + np.copysign(1.0, f0[i, j])
        assert np.array_equal(field[:, j].view(np.uint64), want.view(np.uint64))


def test_the_reference_program_runs_translated_and_prints_the_published_lines(oracle, ref):
    """program biharmonic_wk_scalar_kernel (F90:545-580, GPU1 block restored) with NOTHING but translated reference code between
    the memory and the printed numbers: initialize_data + myrandom, the CPU / GPU1 / GPU2 modules, save_qtens, compute_l2norm.
    The two L2-norm lines are those of atmosphere/results/atm.gnu.4.9.3, atm.gnu.7.1.0 and atm.pgi.17.7, digit for digit.  The
    oracle's own restatements of the harness (LCG fill order, norm) are pinned on the translated ones on the way."""
    prog = ref.Program()
    lines = dict(prog.run())
    assert "%.16E" % lines["GPU1 L2 norm"] == "1.4477731365593603E-16" == "%.16E" % KAT["KAT2_gpu1_vs_cpu_rel_l2"]
    assert "%.16E" % lines["GPU2 L2 norm"] == "1.3341964743086566E-16" == "%.16E" % KAT["KAT1_gpu2_vs_cpu_rel_l2"]
    # initialize_data / compute_l2norm: oracle == translated, bit for bit
    prog.initialize_data()
    dvv, elem, q = oracle.initialize_data(prog.NELEMD, prog.NLEV, prog.QSIZE)
    assert np.array_equal(_bits(dvv), _bits(prog.deriv)) and np.array_equal(_bits(elem), _bits(prog.elem)) and np.array_equal(_bits(q), _bits(prog.qtens))
    prog.save_qtens()
    assert np.array_equal(prog.qtens_save, prog.qtens)
    ref.biharmonic_wk_scalar_gpu2(prog.elem, prog.qtens, prog.deriv)
    assert prog.compute_l2norm() == oracle.compute_l2norm(prog.qtens, prog.qtens_save)
    # a second initialize_data restarts the generator (reset = .true., F90:51 / F90:84-85): same values again
    prog.initialize_data()
    assert np.array_equal(_bits(q), _bits(prog.qtens))
