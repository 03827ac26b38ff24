"""oracle/_ref for the mmf-mpdata-tracer mini-app: the reference's own Fortran subroutines, translated
mechanically by oracle/f90_to_c.py from the source where it lies (this image has no Fortran compiler) and compiled with the
parity flags -- see oracle/ref_mpdata_py.py.  CPU only.

  * the translator itself: semantics of the Fortran subset on a synthetic program (default-real literals are single precision,
    mixed-mode arithmetic, left-to-right max/min, sections, logical IF, bounds checking), refusal of what is outside the subset
  * oracle/mpdata_oracle.c == the translated advect_scalar2D_cpu (mmf-mpdata-tracer/advect_scalar2D_pushncols_openacc.F90:477-642,
    cited A:<line>) bit for bit
  * the reference's own acceptance test (A:679-683: relative L1 error of f and flux of each OpenACC form against the CPU form)
    run on the translations: exactly 0, as the reference's results/advect.gnu.7.1.0 reports
  * the committed SHA-256 pin in tests/golden/mpdata_results.json is the translated reference's output, not only the oracle's
"""
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

DEMO = textwrap.dedent(
    """\
    ! synthetic program for the translator tests (not reference code)
    program demo_prog
      implicit none
      integer, parameter :: n  = 5
      integer, parameter :: rp = selected_real_kind(13)
      integer, parameter :: np1 = n+1
      real(rp) :: scale(n)
      call demo(a, b)
    contains
      subroutine demo(a, b, hits)
        implicit none
        real(rp), intent(in   ) :: a(0:n)
        real(rp), intent(  out) :: b(n, -1:1)
        integer , intent(  out) :: hits(n)
        real(rp) :: eps, t, work(n)
        integer  :: i, k
        logical  :: flag
        real(rp) :: sq
        sq(t) = t*t - 0.1      ! statement function
        flag = .true.
        eps = 1.e-10            ! a default-real literal: float32(1e-10)
        b(:,:) = 0.
        do i = 1, n
          k = max(1, i-1)
          hits(i) = min(n, 7/i, i+3)
          work(i) = 2./(i-k+1)/a(i) + eps + &
                    0.1*a(i-1) + sq(a(i)) * scale(i)
          b(i,-1) = work(i)
          b(i,0) = min(0._rp, a(i), -a(k)) - max(a(i), a(i-1), -0._rp)
          if (flag) then
            if (i >= 2 .and. .not. (i == 4)) b(i,1) = b(i,0) + 1.
          else
            b(i,1) = -1.
          endif
          if (i .le. 2) b(i,1) = b(i,1) - a(0)/3
        enddo
      end subroutine demo
    end program demo_prog
    """
)


def _translate(tmp_path, text, sub, runtime="", check=False, tag="demo"):
    src, out, so = tmp_path / f"{tag}.F90", tmp_path / f"{tag}.c", tmp_path / f"lib{tag}.so"
    src.write_text(text)
    cmd = [sys.executable, os.path.join(ROOT, "oracle", "f90_to_c.py"), str(src), sub, "--out", str(out)]
    if runtime:
        cmd += ["--runtime", runtime]
    if check:
        cmd.append("--check-bounds")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        return None, r.stderr
    subprocess.run(["/usr/bin/gcc", "-O2", "-ffp-contract=off", "-fPIC", "-std=gnu11", "-shared", "-o", str(so), str(out), "-lm"], check=True)
    return C.CDLL(str(so)), out.read_text()


def test_translator_semantics_on_a_synthetic_program(tmp_path):
    L, csrc = _translate(tmp_path, DEMO, "demo", runtime="n")
    assert L is not None, csrc
    assert "1.0e-10f" in csrc and "0.1f" in csrc and "2.0f" in csrc  # default-real literals stay single precision
    rng = np.random.default_rng(7)
    for n in (5, 9):
        a = rng.random(n + 1) - 0.5
        a[2] = 0.0  # a tie / signed-zero case for max and min
        scale = rng.random(n)
        b = np.full((n, 3), np.nan, order="F")
        hits = np.zeros(n, dtype=np.int32)
        L.demo.argtypes = [C.c_int] + [C.c_void_p] * 4
        assert L.demo(n, a.ctypes.data, b.ctypes.data, hits.ctypes.data, scale.ctypes.data) == 0
        f32, f64 = np.float32, np.float64
        want = np.zeros((n, 3))
        with np.errstate(divide="ignore"):  # a(2) = 0: 2./../a(2) is +inf on both sides
            for i in range(1, n + 1):
                k = max(1, i - 1)
                assert hits[i - 1] == min(n, 7 // i, i + 3)
                sq = a[i] * a[i] - f64(f32(0.1))
                w = f64(f32(2.0) / f32(i - k + 1)) / a[i] + f64(f32(1e-10)) + f64(f32(0.1)) * a[i - 1] + sq * scale[i - 1]
                want[i - 1, 0] = w
                lo = 0.0
                for x in (a[i], -a[k]):  # min / max go left to right with < / >
                    lo = x if x < lo else lo
                hi = a[i]
                for x in (a[i - 1], -0.0):
                    hi = x if x > hi else hi
                want[i - 1, 1] = lo - hi
                if i >= 2 and not i == 4:
                    want[i - 1, 2] = want[i - 1, 1] + 1.0
                if i <= 2:
                    want[i - 1, 2] = want[i - 1, 2] - a[0] / 3
        assert np.array_equal(b.view(np.uint64), want.view(np.uint64)), (b, want)


@pytest.mark.parametrize(
    "bad,why",
    [
        ("b(i,0) = a(i)**3", "**"),  # only x**2 (= x*x) is in the subset
        ("do k = 1, n, 2\n b(k,0) = 0.\n enddo", "stride"),
        ("b(i,0) = undeclared_thing + 1.", "undeclared"),
        ("b(i,0) = max(0., 1.)", "default reals"),
        ("b(i,0) = exp(a(i))", "exp"),
        ("b(i,0) = a(i) + b(:,1)", "section"),
        ("goto 10", "outside the subset"),
    ],
)
def test_translator_refuses_what_is_outside_its_subset(tmp_path, bad, why):
    text = DEMO.replace("b(i,-1) = work(i)", bad)
    L, err = _translate(tmp_path, text, "demo", runtime="n", tag="bad")
    assert L is None and "f90_to_c:" in err and why in err, err


def test_translator_bounds_check_skips_and_reports(tmp_path):
    text = DEMO.replace("b(i,-1) = work(i)", "b(i+1,-1) = work(i)")  # writes b(n+1,-1) at i = n
    L, _ = _translate(tmp_path, text, "demo", runtime="n", check=True, tag="oob")
    n = 5
    a, scale = np.linspace(0.1, 1.0, n + 1), np.ones(n)
    guard = np.full(3 * n + 8, -7.0)  # b followed by guard words the skipped statement must not touch
    hits = np.zeros(n, dtype=np.int32)
    L.demo.argtypes = [C.c_int] + [C.c_void_p] * 4
    assert L.demo(n, a.ctypes.data, guard.ctypes.data, hits.ctypes.data, scale.ctypes.data) == 0
    line, dim, sub, upper, name = C.c_int(), C.c_int(), C.c_int(), C.c_int(), C.c_char_p()
    L.demo_violations.restype = C.c_long
    assert L.demo_violations(C.byref(line), C.byref(dim), C.byref(sub), C.byref(upper), C.byref(name)) == 1
    assert (dim.value, sub.value, upper.value, name.value) == (1, n + 1, n, b"b")
    assert DEMO.splitlines()[line.value - 1].strip().startswith("b(i,-1)")
    assert np.all(guard[3 * n:] == -7.0) and guard[0] == 0.0 and guard[1] != 0.0


# ------------------------------------------------------------------------------------------------ the reference itself
@pytest.fixture(scope="module")
def ref():
    import ref_mpdata_py

    if not ref_mpdata_py.available():
        pytest.skip("oracle/_ref/libmpdata_f90ref.so is absent and /root/reference is not here to build it")
    try:
        ref_mpdata_py.lib()
    except OSError as e:  # the checker itself could not be loaded on this box: nothing was compared
        pytest.skip(f"oracle/_ref library did not load: {e}")
    return ref_mpdata_py


@pytest.fixture(scope="module")
def mpdata():
    import __graft_entry__ as ge

    ge.build()
    from codesign_kernels_b200 import mpdata as m

    return m


@pytest.fixture(scope="module")
def mpo():
    import mpdata_oracle_py

    return mpdata_oracle_py


@pytest.mark.parametrize(
    "ns,nx,nz,signed,seed",
    [(48, 32, 58, False, 100), (48, 32, 58, True, 2), (5, 7, 9, True, 3), (3, 1, 3, True, 4), (100, 60, 40, True, 5), (1, 32, 58, True, 6),
     (300, 32, 58, True, 7)],
)
def test_oracle_is_bit_identical_to_the_translated_reference(mpdata, mpo, ref, ns, nx, nz, signed, seed):
    """oracle/mpdata_oracle.c against the reference's advect_scalar2D_cpu (A:477-642) at the reference sizes (A:6-8) and at
    others (the source marks nslices / nz / nx "User changeable"), unsigned velocities as the reference draws them (A:654-660)
    and signed ones that drive both upwind branches."""
    P = mpdata.TestAdvect(ns, nx, nz).init(seed=seed, signed_velocity=signed)
    f0, fl0 = P.f.copy(order="F"), P.flux.copy(order="F")
    mpo.advect_scalar2D_cpu(ns, nx, nz, f0, P.u, P.w, P.rho, P.rhow, P.adz, fl0)
    f1, fl1 = P.f.copy(order="F"), P.flux.copy(order="F")
    ref.advect_scalar2D_cpu(ns, nx, nz, f1, P.u, P.w, P.rho, P.rhow, P.adz, fl1)
    assert np.isfinite(f1).all() and np.isfinite(fl1[:, :-1]).all()  # no NaN-poisoned local of the translation reaches an output
    assert np.array_equal(f0.view(np.uint64), f1.view(np.uint64)) and np.array_equal(fl0.view(np.uint64), fl1.view(np.uint64))
    assert np.array_equal(fl1[:, -1], P.flux[:, -1])  # flux(:,nz) is never written (A:540-541 run k = 1..nzm)


def test_reference_acceptance_test_on_the_translations(mpdata, ref):
    """program test_advect's own check (A:49-66 with the commented calls restored, compare() A:679-683): each OpenACC form
    against the CPU form, relative L1 error of f and of flux.  Without reassociation or FMA contraction both are exactly 0 --
    what the reference's results/advect.gnu.7.1.0 reports for its CPU builds of the annotated forms."""
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "mpdata_results.json")))
    gnu = gold["reference_results"]["advect.gnu.7.1.0"]
    assert set(gnu["rel_l1_f"]) == {0.0} and set(gnu["rel_l1_flux"]) == {0.0}
    for signed in (False, True):
        P = mpdata.TestAdvect().init(seed=100, signed_velocity=signed)  # 48 x 32 x 58, A:6-8
        out = {}
        for form in ref.FORMS:
            f, fl = P.f.copy(order="F"), P.flux.copy(order="F")
            ref.advect_scalar2D(form, P.nslices, P.nx, P.nz, f, P.u, P.w, P.rho, P.rhow, P.adz, fl)
            out[form] = (f, fl)
        for form in ("openacc_1", "openacc_2"):
            rel_f = float(np.abs(out[form][0] - out["cpu"][0]).sum() / np.abs(out["cpu"][0]).sum())
            rel_fl = float(np.abs(out[form][1] - out["cpu"][1]).sum() / np.abs(out["cpu"][1]).sum())
            assert rel_f == 0.0 and rel_fl == 0.0, (form, rel_f, rel_fl)
        # a finding about the reference: A:126 `if (k == 1) www(sl,i,j,nz) = 0.` runs for i = -1..nxp3 but www is declared
        # -1:nxp2 (A:83): one write per slice lands past the array.  The checked build skips and counts them.
        n, line, name, dim, sub, upper = ref.violations()
        assert (n, line, name, dim, sub, upper) == (P.nslices, 126, "www", 2, P.nx + 3, P.nx + 2)


def test_committed_pin_is_the_translated_references_output(mpdata, ref):
    """tests/golden/mpdata_results.json: the SHA-256 of f and flux after one step on the integer-LCG inputs was written as a
    regression pin of the oracle; the translated reference gives the same bytes, so the GPU test that compares the CUDA result
    with that pin (test_gpu_matches_committed_fixture) compares it with the reference's arithmetic."""
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_mpdata_golden as mk

    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "mpdata_results.json")))
    a = mk.lcg_fields()
    f, fl = a["f"].copy(order="F"), a["flux"].copy(order="F")
    ref.advect_scalar2D_cpu(48, 32, 58, f, a["u"], a["w"], a["rho"], a["rhow"], a["adz"], fl)
    got = {"f_sha256": hashlib.sha256(f.tobytes(order="F")).hexdigest(), "flux_sha256": hashlib.sha256(fl.tobytes(order="F")).hexdigest()}
    pin = gold["translated_reference_pin"]
    assert got["f_sha256"] == pin["f_sha256"] and got["flux_sha256"] == pin["flux_sha256"]
    assert pin["f_sha256"] == gold["oracle_regression_pin"]["f_sha256"] and pin["flux_sha256"] == gold["oracle_regression_pin"]["flux_sha256"]
