import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import __graft_entry__ as ge
bwk = ge.build()
import oracle_py as orc
reps = int(sys.argv[1])
cfgs = [tuple(int(x) for x in a.split(",")) for a in sys.argv[2:]]
refs = {}
info = {}
for (ne, nlev, qsize) in cfgs:
    n = 6 * ne * ne
    dvv, elem, q = orc.initialize_data(n, nlev, qsize)
    dss = orc.Dss(n, orc.cubed_sphere_gdof(ne))
    rs = dss.rspheremp(elem)
    refs[(ne, nlev, qsize)] = dss.biharmonic_full(elem, rs, q.copy(order="F"), dvv, omp=True)
    pl = bwk.Plan(n, bwk.cubed_sphere_gdof(ne))
    f, c = pl.strips(0); reg = pl.strip_regular
    regular_of = np.zeros(n, int); pos_of = np.zeros(n, int); cnt_of = np.zeros(n, int)
    for a, k, r in zip(f, c, reg):
        regular_of[a:a+k] = r; pos_of[a:a+k] = np.arange(k); cnt_of[a:a+k] = k
    info[(ne, nlev, qsize)] = (regular_of, pos_of, cnt_of)
nfail = 0
for rep in range(reps):
    for cfg in cfgs:
        ne, nlev, qsize = cfg
        n = 6 * ne * ne
        with bwk.KernelUtils(n, nlev, qsize).allocate_data(host_arrays=False) as ku:
            ku.initialize_data_device()
            ku.set_topology_cubed_sphere(ne)
            ku.run_biharmonic_device()
            got = ku.download_elements(0, n)
        ref = refs[cfg]
        rel = np.abs(got - ref) / np.abs(ref)
        if rel.max() > 1e-12:
            nfail += 1
            idx = np.nonzero(rel > 1e-12)
            be = np.unique(idx[4])
            kq = np.unique((idx[2] + nlev * idx[3]) // 32)
            regular_of, pos_of, cnt_of = info[cfg]
            print(f"rep {rep} cfg {cfg}: maxrel {rel.max():.2e} bad elems {be[:8]} regular {regular_of[be][:8]} pos {pos_of[be][:8]} cnt {cnt_of[be][:8]} tiles {kq[:8]}", flush=True)
print(f"{nfail} failing runs of {reps * len(cfgs)}")
