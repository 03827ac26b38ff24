import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
bwk = ge.build(oracle=False)
ne, nlev, qsize = 2, 8, 4
n = 6 * ne * ne
ku = bwk.KernelUtils(n, nlev, qsize).allocate_data(host_arrays=False)
ku.initialize_data_device()
ku.set_topology_cubed_sphere(ne)
ku.get_rspheremp()
ku.run_biharmonic_device()
ku.sync()
print("ok")
