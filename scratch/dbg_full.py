import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import __graft_entry__ as ge
bwk = ge.build()
import oracle_py as orc

def run(ne, nlev, qsize):
    n = 6 * ne * ne
    dvv, elem, q = orc.initialize_data(n, nlev, qsize)
    dss = orc.Dss(n, orc.cubed_sphere_gdof(ne))
    rs = dss.rspheremp(elem)
    ref = dss.biharmonic_full(elem, rs, q.copy(order="F"), dvv)
    with bwk.KernelUtils(n, nlev, qsize).allocate_data(host_arrays=False) as ku:
        ku.initialize_data_device()
        ku.set_topology_cubed_sphere(ne)
        ku.run_biharmonic_device()
        got = ku.download_elements(0, n)
    rel = np.abs(got - ref) / np.abs(ref)
    bad = rel > 1e-12
    print(f"ne{ne} nlev{nlev} q{qsize}: maxrel {rel.max():.3e} bad {bad.sum()} of {bad.size}")
    if bad.any():
        pl = bwk.Plan(n, bwk.cubed_sphere_gdof(ne))
        f, c = pl.strips(0); reg = pl.strip_regular
        regular_of = np.zeros(n, int); pos_of = np.zeros(n, int); cnt_of = np.zeros(n, int)
        for a, k, r in zip(f, c, reg):
            regular_of[a:a+k] = r; pos_of[a:a+k] = np.arange(k); cnt_of[a:a+k] = k
        be = np.unique(np.nonzero(bad)[4])
        print(" bad elements:", len(be), "regular:", regular_of[be].sum(), "first few", be[:10])
        print(" pos hist", np.bincount(pos_of[be], minlength=4), "cnt hist", np.bincount(cnt_of[be], minlength=5))
        bk = np.unique(np.nonzero(bad)[2] + nlev * np.nonzero(bad)[3])
        print(" bad kq:", bk[:40], "count", len(bk))
        # bad points within slab (errors propagate through the laplacian, so all 16 are usually bad)
        e0 = be[0]
        print(" elem", e0, "rel err slab kq0:", rel[:, :, bk[0] % nlev, bk[0] // nlev, e0].round(3))

for a in sys.argv[1:]:
    run(*[int(x) for x in a.split(",")])
