import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
dbg = int(sys.argv[1])
os.environ["BWK_DBG"] = str(dbg)
import __graft_entry__ as ge
bwk = ge.build()
import oracle_py as orc
ne, nlev, qsize = 30, 72, 1
n = 6 * ne * ne
dvv, elem, q = orc.initialize_data(n, nlev, qsize)
dss = orc.Dss(n, orc.cubed_sphere_gdof(ne))
rs = dss.rspheremp(elem)
lap = orc.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv)
ref = np.empty_like(lap)
for k in range(nlev):
    ref[:, :, k:k+1, :, :] = dss.apply(np.asfortranarray(lap[:, :, k:k+1, :, :]))
rs4 = np.asarray(rs).reshape(4, 4, 1, 1, n, order="F")
if dbg == 16:
    ref = ref * rs4
if dbg == 32:
    ref = orc.biharmonic_wk_scalar_cpu(elem, np.asfortranarray(ref), dvv)
for rep in range(2):
    with bwk.KernelUtils(n, nlev, qsize).allocate_data(host_arrays=False) as ku:
        ku.initialize_data_device()
        ku.set_topology_cubed_sphere(ne)
        if rep == 0:
            print("rsph equal to oracle:", np.array_equal(np.asarray(ku.get_rspheremp()).reshape(-1), np.asarray(rs).reshape(-1)))
        ku.run_biharmonic_device()
        got = ku.download_elements(0, n)
    rel = np.abs(got - ref) / np.abs(ref)
    bad = rel > 1e-12
    print(f"dbg {dbg} rep {rep}: maxrel {rel.max():.3e} bad {bad.sum()}")
    if bad.any():
        idx = np.nonzero(bad)
        be = np.unique(idx[4]); print(" bad elems", be[:10])
        e0 = be[0]
        ks = np.unique(idx[2][idx[4] == e0]); print(" bad k", ks)
        k0 = ks[0]
        print(" bad mask (i rows, j cols):\n", bad[:, :, k0, 0, e0].astype(int))
        print(" got/ref\n", got[:, :, k0, 0, e0] / ref[:, :, k0, 0, e0])
        print(" rs\n", rs4[:, :, 0, 0, e0])
        print(" spheremp\n", elem[e0][64:80] if hasattr(elem, '__getitem__') else None)
