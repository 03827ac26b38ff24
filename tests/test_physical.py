"""Analytic known-answer tests with PHYSICAL metric terms: true GLL derivative matrix
and gnomonic cubed-sphere Dinv / spheremp, identity tensorVisc.  On these inputs the operator of
atmosphere/biharmonic_wk_kernel.F90:109-182 is HOMME's weak spherical Laplacian, so spherical harmonics are
eigenfunctions:
  * pointwise:  DSS(laplace_sphere_wk(f)) * rspheremp  ->  -l(l+1) rrearth^2 f       (error O(h^3) for np = 4)
  * energy:     sum_e sum_pts f * [Laplace -> DSS -> rspheremp -> Laplace](f)  =  ||discrete lap f||^2
                ->  (l(l+1) rrearth^2)^2 * sum spheremp f^2                          (error O(h^6))
The second one pins the semantics of the DSS step + second sweep (which the reference does not contain, F90:518-519)
to an analytic value instead of to the builder's own restatement.  CPU tests run the C oracle; the GPU tests run
libbwk_b200 through the C ABI on the same inputs.
"""
import numpy as np
import pytest

NLEV, QSIZE = 3, 2  # six slots -> harmonics of degree 1..6


def _cs():
    from codesign_kernels_b200 import cubed_sphere

    return cubed_sphere


def _setup(oracle, ne):
    cs = _cs()
    dvv, elem, xyz = cs.physical_elements(ne)
    n = 6 * ne * ne
    dss = oracle.Dss(n, oracle.cubed_sphere_gdof(ne))
    rs = dss.rspheremp(elem)
    q0, deg = cs.harmonic_field(xyz, NLEV, QSIZE)
    mp = elem[:, 64:80].reshape(n, 4, 4).transpose(2, 1, 0)  # spheremp as [i][j][e]
    return dvv, elem, dss, rs, q0, deg, mp


def _pointwise_errors(oracle, lap_weak, dss, rs, q0, deg):
    """relative L2 error of DSS(weak laplacian)*rspheremp against -l(l+1) rrearth^2 f, per slot"""
    rr = oracle.rrearth()
    u = lap_weak.copy(order="F")
    dss.apply(u)
    u *= rs[:, :, None, None, :]
    out = {}
    for t in range(QSIZE):
        for k in range(NLEV):
            lam = deg[k, t] * (deg[k, t] + 1) * rr * rr
            f = q0[:, :, k, t, :]
            out[int(deg[k, t])] = float(np.sqrt(((u[:, :, k, t, :] + lam * f) ** 2).sum() / ((lam * f) ** 2).sum()))
    return out


def _energy_errors(oracle, bih_weak, q0, deg, mp):
    rr = oracle.rrearth()
    out = {}
    for t in range(QSIZE):
        for k in range(NLEV):
            lam = deg[k, t] * (deg[k, t] + 1) * rr * rr
            f = q0[:, :, k, t, :]
            out[int(deg[k, t])] = float((f * bih_weak[:, :, k, t, :]).sum() / (lam * lam * (mp * f * f).sum()) - 1.0)
    return out


def test_gll4_facts():
    x, w, d = _cs().gll4()
    assert abs(w.sum() - 2.0) < 1e-15 and np.allclose(x, -x[::-1])
    assert np.abs(d.sum(axis=1)).max() < 1e-14  # derivative of a constant
    for p in range(1, 4):  # exact for polynomials up to degree np-1
        assert np.allclose(d @ x**p, p * x ** (p - 1), atol=1e-13)
    # quadrature exact up to degree 2 np - 3 = 5
    assert abs((w * x**4).sum() - 2.0 / 5.0) < 1e-15


@pytest.mark.parametrize("ne", [2, 4, 8])
def test_sphere_area_and_topology_match(oracle, ne):
    """sum spheremp -> 4 pi, and the geometric points coincide exactly where the library's synthetic topology says so"""
    cs = _cs()
    _, elem, xyz = cs.physical_elements(ne)
    assert abs(elem[:, 64:80].sum() / (4 * np.pi) - 1.0) < {2: 1e-5, 4: 2e-7, 8: 4e-9}[ne]
    g = oracle.cubed_sphere_gdof(ne).reshape(-1, 4, 4)  # [e][j][i]
    order = np.argsort(g, axis=None, kind="stable")
    gs = g.reshape(-1)[order]
    ps = xyz.reshape(-1, 3)[order]
    same = gs[1:] == gs[:-1]
    assert np.abs(ps[1:][same] - ps[:-1][same]).max() < 1e-14          # shared nodes are the same point of the sphere
    assert np.linalg.norm(ps[1:][~same] - ps[:-1][~same], axis=1).min() > 1e-3 / ne  # distinct nodes are distinct points


def test_laplacian_of_harmonics_converges_cpu(oracle):
    errs = {}
    for ne in (4, 8):
        dvv, elem, dss, rs, q0, deg, mp = _setup(oracle, ne)
        lap = oracle.biharmonic_wk_scalar_cpu(elem, q0.copy(order="F"), dvv.reshape(16))
        errs[ne] = _pointwise_errors(oracle, lap, dss, rs, q0, deg)
    for l in range(1, 7):
        assert errs[8][l] < 5e-3 and errs[8][l] < errs[4][l] / 5.0, (l, errs[4][l], errs[8][l])  # ~ h^3
    assert max(errs[8][l] for l in (1, 2, 3, 4)) < 1.6e-3


def test_full_biharmonic_energy_known_answer_cpu(oracle):
    errs = {}
    for ne in (4, 8):
        dvv, elem, dss, rs, q0, deg, mp = _setup(oracle, ne)
        b = dss.biharmonic_full(elem, rs, q0.copy(order="F"), dvv.reshape(16))
        errs[ne] = _energy_errors(oracle, b, q0, deg, mp)
    for l in range(1, 7):
        assert abs(errs[8][l]) < 2.5e-5 and abs(errs[8][l]) < abs(errs[4][l]) / 15.0, (l, errs[4][l], errs[8][l])  # ~ h^6
    assert max(abs(errs[8][l]) for l in (1, 2, 3)) < 5e-7


@pytest.mark.gpu
@pytest.mark.parametrize("ne", [4, 8])
def test_physical_terms_gpu(bwk, oracle, ne):
    """the CUDA path on physical inputs: analytic values, and agreement with the oracle.  Dvv has mixed signs here,
    so the contractions cancel and the elementwise-relative bound of the LCG tests is replaced by a bound relative
    to the largest magnitude of each (k,q) field: 1e-12 for one Laplacian (measured 3e-15 / 8e-15 at ne 4 / 8).  The
    second Laplacian differentiates the rounding noise of the first (amplification ~ lambda_max / lambda_l of the
    discrete operator, growing like ne^2), so the bound for the full biharmonic is derived from the oracle itself: ten
    times the change of the oracle's result when its own sweep-1 output is perturbed by 1e-14 of the field maximum
    (measured GPU-vs-oracle: 1.0e-12 at ne=4, 1.1e-11 at ne=8)."""
    dvv, elem, dss, rs, q0, deg, mp = _setup(oracle, ne)
    n = 6 * ne * ne
    ref_lap = oracle.biharmonic_wk_scalar_cpu(elem, q0.copy(order="F"), dvv.reshape(16))
    ref_bih = dss.biharmonic_full(elem, rs, q0.copy(order="F"), dvv.reshape(16))
    with bwk.KernelUtils(n, NLEV, QSIZE).allocate_data() as ku:
        ku.elem.view(np.float64).reshape(n, 144)[...] = elem
        ku.deriv["Dvv"] = dvv
        ku.qtens[...] = q0
        ku.set_topology_cubed_sphere(ne)
        ku.biharmonic_wk_scalar(ku.elem, ku.qtens, None, ku.deriv, 1, n)
        lap = ku.qtens.copy(order="F")
        ku.qtens[...] = q0
        ku.biharmonic_wk_scalar_full(ku.elem, ku.qtens, ku.deriv)
        bih = ku.qtens.copy(order="F")
        assert np.allclose(ku.get_rspheremp(), rs, rtol=1e-15, atol=0)

    def slot_rel(a, b):
        worst = 0.0
        for t in range(QSIZE):
            for k in range(NLEV):
                worst = max(worst, float(np.abs(a[:, :, k, t, :] - b[:, :, k, t, :]).max() / np.abs(b[:, :, k, t, :]).max()))
        return worst

    assert slot_rel(lap, ref_lap) <= 1e-12, slot_rel(lap, ref_lap)
    rng = np.random.default_rng(5)
    pert = ref_lap.copy(order="F")
    for t in range(QSIZE):
        for k in range(NLEV):
            pert[:, :, k, t, :] += 1e-14 * np.abs(ref_lap[:, :, k, t, :]).max() * rng.choice([-1.0, 1.0], size=(4, 4, n))
    dss.apply(pert)
    pert *= rs[:, :, None, None, :]
    oracle.biharmonic_wk_scalar_cpu(elem, pert, dvv.reshape(16))
    bound = 10.0 * slot_rel(pert, ref_bih) + 1e-12
    print(f"ne={ne}: GPU vs oracle, relative to max: laplace {slot_rel(lap, ref_lap):.2e}, biharmonic {slot_rel(bih, ref_bih):.2e} (bound {bound:.2e})")
    assert slot_rel(bih, ref_bih) <= bound, (slot_rel(bih, ref_bih), bound)
    pe = _pointwise_errors(oracle, lap, dss, rs, q0, deg)
    ee = _energy_errors(oracle, bih, q0, deg, mp)
    if ne == 8:
        assert max(pe[l] for l in (1, 2, 3, 4)) < 1.6e-3 and max(pe.values()) < 5e-3, pe
        assert max(abs(v) for v in ee.values()) < 2.5e-5, ee
    else:
        assert max(pe.values()) < 4e-2 and max(abs(v) for v in ee.values()) < 1.2e-3, (pe, ee)
