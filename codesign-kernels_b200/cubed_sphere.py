"""Physical element data for the biharmonic_wk path: true GLL derivative matrix and gnomonic
cubed-sphere metric terms, as a host-side (numpy) generator.

The reference driver fills `deriv%Dvv` and `elem(:)%Dinv/spheremp/tensorVisc` with LCG noise
(atmosphere/biharmonic_wk_kernel.F90:48-58), which exercises the arithmetic but has no analytic answer.
With the terms below the same operator is the weak spherical Laplacian of HOMME, so

    DSS(laplace_sphere_wk(f)) * rspheremp  ~=  -l(l+1) rrearth^2 f        for a spherical harmonic f of degree l,

and the full biharmonic gives (l(l+1))^2 rrearth^4 f -- a known-answer test of the operator AND of the DSS
step, which the reference does not contain (only the comment F90:518-519).

Conventions (they follow how the reference consumes the terms):
  * Dvv(i,l) = d(basis i)/dxi at GLL point l            (gradient_sphere, F90:116-127: v1(l,j) = sum_i Dvv(i,l) s(i,j))
  * D[r][c] = e_r . dP/dxi_c, P on the unit sphere, (e_1, e_2) an orthonormal tangent frame; Dinv(i,j,a,b) = (D^-1)[a][b]
    (F90:128-133 applies Dinv^T to the reference-space gradient, F90:145-150 applies Dinv to the vector)
  * spheremp(i,j) = w_i w_j |det D|                     (F90:151-159)
  * tensorVisc(i,j,:,:) = identity                      (plain Laplacian, F90:172-180)
  * the radius enters only through rrearth (F90:14), as in the reference.
Element order and face maps are those of the library's synthetic cubed sphere (bwk_set_topology_cubed_sphere,
SURVEY.md App. D): e = (f*ne + je)*ne + ie, faces f0:(1,X,Y) f1:(-X,1,Y) f2:(-1,-X,Y) f3:(X,-1,Y) f4:(X,Y,1) f5:(X,Y,-1),
equiangular gnomonic coordinates X = tan(alpha), Y = tan(beta).
"""
import numpy as np

NP = 4


def gll4():
    """np=4 Gauss-Lobatto-Legendre nodes, weights and differentiation matrix D[l][i] = basis_i'(xi_l)."""
    s5 = np.sqrt(5.0)
    x = np.array([-1.0, -1.0 / s5, 1.0 / s5, 1.0])
    w = np.array([1.0, 5.0, 5.0, 1.0]) / 6.0
    d = np.zeros((NP, NP))
    for i in range(NP):
        others = [m for m in range(NP) if m != i]
        denom = np.prod([x[i] - x[m] for m in others])
        for l in range(NP):
            acc = 0.0
            for m in others:
                acc += np.prod([x[l] - x[r] for r in others if r != m])
            d[l, i] = acc / denom
    return x, w, d


def dvv_array():
    """deriv%Dvv as the (4,4) array of DERIVATIVE_DTYPE / the oracle's flat dvv: element [l][i] = Dvv_f(i,l)."""
    return gll4()[2].copy()


_FACE = (
    # C(X,Y), dC/dX, dC/dY
    (lambda X, Y: (1 + 0 * X, X, Y), (0, 1, 0), (0, 0, 1)),
    (lambda X, Y: (-X, 1 + 0 * X, Y), (-1, 0, 0), (0, 0, 1)),
    (lambda X, Y: (-1 + 0 * X, -X, Y), (0, -1, 0), (0, 0, 1)),
    (lambda X, Y: (X, -1 + 0 * X, Y), (1, 0, 0), (0, 0, 1)),
    (lambda X, Y: (X, Y, 1 + 0 * X), (1, 0, 0), (0, 1, 0)),
    (lambda X, Y: (X, Y, -1 + 0 * X), (1, 0, 0), (0, 1, 0)),
)


def physical_elements(ne, axis=None, with_frame=False):
    """-> (dvv (4,4), elem (6 ne^2, 144) float64 records in element_t order (F90:23-27), xyz (6 ne^2, 4[j], 4[i], 3)).

    axis=None: the tangent frame is local to each face (e_1 along dP/dxi_1) -- enough for scalar fields, whose operator does
    not depend on the frame.  axis=(ax,ay,az): a GLOBAL frame, e_1 = axis x P (east about that axis), e_2 = P x e_1 (north),
    the same at every copy of a shared node, so that vector components can be summed by the DSS (HOMME's lat-lon
    components; the axis is tilted off the cube's symmetry axes so that its poles are not GLL nodes).
    with_frame=True appends (vec_sphere2cart (6 ne^2, 96) in Fortran order (np,np,3,2), e1, e2 (6 ne^2, 4, 4, 3))."""
    x, w, d = gll4()
    if axis is not None:
        axis = np.asarray(axis, float) / np.linalg.norm(axis)
    n = 6 * ne * ne
    elem = np.zeros((n, 144))
    xyz = np.zeros((n, NP, NP, 3))
    v2c = np.zeros((n, 96))
    f1 = np.zeros((n, NP, NP, 3))
    f2 = np.zeros((n, NP, NP, 3))
    dalpha = np.pi / (4.0 * ne)  # d(alpha)/d(xi)
    for f in range(6):
        cfun, cu, cv = _FACE[f]
        cu, cv = np.array(cu, float), np.array(cv, float)
        for je in range(ne):
            for ie in range(ne):
                e = (f * ne + je) * ne + ie
                alpha = -np.pi / 4 + (np.pi / 2) * (ie + (x + 1) / 2) / ne  # [i]
                beta = -np.pi / 4 + (np.pi / 2) * (je + (x + 1) / 2) / ne   # [j]
                X = np.tan(alpha)[None, :] + np.zeros((NP, 1))              # [j][i]
                Y = np.tan(beta)[:, None] + np.zeros((1, NP))
                C = np.stack(cfun(X, Y), axis=-1)
                r = np.linalg.norm(C, axis=-1, keepdims=True)
                P = C / r
                xyz[e] = P
                # covariant basis a_c = dP/dxi_c (tangent to the unit sphere)
                a1 = (cu - P * (P @ cu)[..., None]) / r * ((1 + X * X) * dalpha)[..., None]
                a2 = (cv - P * (P @ cv)[..., None]) / r * ((1 + Y * Y) * dalpha)[..., None]
                if axis is None:
                    e1 = a1 / np.linalg.norm(a1, axis=-1, keepdims=True)
                else:
                    e1 = np.cross(axis, P)
                    e1 = e1 / np.linalg.norm(e1, axis=-1, keepdims=True)
                e2 = np.cross(P, e1)
                f1[e], f2[e] = e1, e2
                for c in range(3):  # vec_sphere2cart(i,j,c,m) at i + 4j + 16c + 48m
                    v2c[e, 16 * c:16 * c + 16] = e1[..., c].reshape(16)
                    v2c[e, 48 + 16 * c:48 + 16 * c + 16] = e2[..., c].reshape(16)
                D = np.empty((NP, NP, 2, 2))
                D[..., 0, 0] = np.sum(e1 * a1, -1)
                D[..., 0, 1] = np.sum(e1 * a2, -1)
                D[..., 1, 0] = np.sum(e2 * a1, -1)
                D[..., 1, 1] = np.sum(e2 * a2, -1)
                det = D[..., 0, 0] * D[..., 1, 1] - D[..., 0, 1] * D[..., 1, 0]
                Di = np.empty_like(D)
                Di[..., 0, 0] = D[..., 1, 1] / det
                Di[..., 0, 1] = -D[..., 0, 1] / det
                Di[..., 1, 0] = -D[..., 1, 0] / det
                Di[..., 1, 1] = D[..., 0, 0] / det
                rec = elem[e]
                # Fortran Dinv(i,j,a,b) at i + 4j + 16a + 32b ; arrays here are indexed [j][i]
                for a in range(2):
                    for b in range(2):
                        rec[16 * a + 32 * b:16 * a + 32 * b + 16] = Di[..., a, b].reshape(16)
                        rec[80 + 16 * a + 32 * b:80 + 16 * a + 32 * b + 16] = 1.0 if a == b else 0.0
                rec[64:80] = (np.abs(det) * w[None, :] * w[:, None]).reshape(16)
    if with_frame:
        return d.copy(), elem, xyz, v2c, f1, f2
    return d.copy(), elem, xyz


TILTED_AXIS = (0.31, 0.52, 0.79)


def harmonic_vector(xyz, l, e1, e2, rotational=False):
    """(u, v) components, in the frame (e1, e2), of the tangent gradient of harmonic(l) -- or of its rotation by 90 degrees
    (P x grad).  Both are eigenfunctions of the vector Laplacian (grad div - curl curl) with eigenvalue -l(l+1); the
    rotational l = 1 field is a rigid rotation.  -> array (..., 2)"""
    z = (xyz[..., 0] + 1j * xyz[..., 1]) ** (l - 1)
    g = np.stack([l * z.real, -l * z.imag, 0 * z.real], axis=-1)
    g = g - xyz * np.sum(xyz * g, axis=-1, keepdims=True)
    if rotational:
        g = np.cross(xyz, g)
    return np.stack([np.sum(g * e1, -1), np.sum(g * e2, -1)], axis=-1)


def harmonic(xyz, l):
    """Re (x + i y)^l on the unit sphere: a spherical harmonic of degree l (Laplacian eigenvalue -l(l+1))."""
    z = (xyz[..., 0] + 1j * xyz[..., 1]) ** l
    return z.real


def harmonic_field(xyz, nlev, qsize, degrees=None):
    """qtens (4,4,nlev,qsize,nelem) Fortran order with harmonic(l) in slot (k,q), l = degrees[k + nlev*q] (default 1,2,3,...).
    -> (qtens, degree per slot (nlev, qsize))."""
    n = xyz.shape[0]
    q = np.zeros((NP, NP, nlev, qsize, n), order="F")
    deg = np.zeros((nlev, qsize), dtype=np.int64)
    for t in range(qsize):
        for k in range(nlev):
            l = (degrees[k + nlev * t] if degrees is not None else 1 + (k + nlev * t) % 6)
            deg[k, t] = l
            q[:, :, k, t, :] = np.transpose(harmonic(xyz, l), (2, 1, 0))  # [e][j][i] -> [i][j][e]
    return q, deg
