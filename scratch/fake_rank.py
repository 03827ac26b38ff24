"""profiling aid: the launches of rank r of n on one GPU (BWK_FAKE_RANKS=r,n BWK_TRACE=1), ne120 L72 q40"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
bwk = ge.build(oracle=False)
r, nr = (int(x) for x in os.environ["BWK_FAKE_RANKS"].split(","))
ne, nlev, qsize = 120, 72, 40
n = 6 * ne * ne
first, count = bwk.partition(n, r, nr)
ku = bwk.KernelUtils(count, nlev, qsize).allocate_data(host_arrays=False)
ku.initialize_data_device(n, first)
ku.set_topology_cubed_sphere(ne, first)
ku.get_rspheremp()
for rep in range(5):
    ku.initialize_data_device(n, first)
    ku.run_biharmonic_device()
    ku.sync()
ku.finalize()
