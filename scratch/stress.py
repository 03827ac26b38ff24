import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import __graft_entry__ as ge
bwk = ge.build()
import oracle_py as orc
ne, nlev, qsize, reps = (int(x) for x in sys.argv[1:5])
n = 6 * ne * ne
dvv, elem, q = orc.initialize_data(n, nlev, qsize)
dss = orc.Dss(n, orc.cubed_sphere_gdof(ne))
rs = dss.rspheremp(elem)
ref = dss.biharmonic_full(elem, rs, q.copy(order="F"), dvv, omp=True)
pl = bwk.Plan(n, bwk.cubed_sphere_gdof(ne))
f, c = pl.strips(0); reg = pl.strip_regular
regular_of = np.zeros(n, int); pos_of = np.zeros(n, int); cnt_of = np.zeros(n, int)
for a, k, r in zip(f, c, reg):
    regular_of[a:a+k] = r; pos_of[a:a+k] = np.arange(k); cnt_of[a:a+k] = k
nfail = 0
first = None
with bwk.KernelUtils(n, nlev, qsize).allocate_data(host_arrays=False) as ku:
    ku.initialize_data_device()
    ku.set_topology_cubed_sphere(ne)
    for rep in range(reps):
        ku.initialize_data_device()
        ku.run_biharmonic_device()
        got = ku.download_elements(0, n)
        if first is None:
            first = got
            rel = np.abs(got - ref) / np.abs(ref)
            print("first rep maxrel vs oracle", rel.max())
        if not np.array_equal(got, first) or rep == 0 and rel.max() > 1e-12:
            nfail += 1
            bad = got != first if rep else rel > 1e-12
            idx = np.nonzero(bad)
            be = np.unique(idx[4])
            kq = np.unique((idx[2] + nlev * idx[3]) // 32)
            print(f"rep {rep}: bad elems {be[:8]} regular {regular_of[be][:8]} pos {pos_of[be][:8]} cnt {cnt_of[be][:8]} tiles {kq[:8]}")
print(f"ne{ne} L{nlev} q{qsize}: {nfail} failing reps of {reps}")
