"""TEST INFRASTRUCTURE ONLY -- the "patch oracle": the full biharmonic of ONE element of a big cubed sphere, assembled
from the CPU oracle's pieces on the element's own neighbourhood, so that BASELINE's 31.85 GB / 257.7 GB configurations
can be checked against the oracle at chosen elements without a whole-sphere CPU run.  Bit-identical to the full oracle
(oracle.Dss.biharmonic_full) at small ne (tests/test_plan.py::test_patch_oracle_equals_full_oracle)."""
import numpy as np


def patch_oracle_full(oracle, ne, nlev, qsize, gd, e):
    """Full biharmonic of ONE element e of a big cubed sphere from the oracle's pieces: the CPU-order Laplacian
    (F90:109-200) of e and of every element sharing a node with it, the DSS sum over coincident points in ascending
    global point order, rspheremp from the same sums of spheremp, then the Laplacian of e again."""
    n = 6 * ne * ne
    keys = gd[e]
    nbrs = np.unique(np.nonzero(np.isin(gd, keys))[0])
    lap, sph = {}, {}
    for m in nbrs:
        dvv, elem, q = oracle.initialize_subset(n, nlev, qsize, int(m), 1)
        sph[int(m)] = elem[0, 64:80].copy()
        lap[int(m)] = oracle.biharmonic_wk_scalar_cpu(elem, q, dvv).reshape(16, nlev * qsize, order="F")
        if m == e:
            dvv_e, elem_e = dvv, elem
    out = np.empty((16, nlev * qsize))
    for pt in range(16):
        members = [(int(m), int(t)) for m in nbrs for t in np.nonzero(gd[m] == keys[pt])[0]]
        members.sort()
        s = None
        w = None
        for m, t in members:
            s = lap[m][t].copy() if s is None else s + lap[m][t]
            w = sph[m][t] if w is None else w + sph[m][t]
        out[pt] = s  # the oracle scales by rspheremp inside its second sweep (element_sweep2)
        out[pt] *= 1.0 / w
    q2 = np.asfortranarray(out.reshape((4, 4, nlev, qsize, 1), order="F"))
    return oracle.biharmonic_wk_scalar_cpu(elem_e, q2, dvv_e)
