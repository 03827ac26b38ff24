"""One rank's launch lists of an N-rank run on ONE GPU (BWK_FAKE_RANKS=r,n with BWK_TRACE=1: phases serialised and timed, halo rows
never filled, results meaningless): what a rank's kernels cost without any communication.
usage: BWK_FAKE_RANKS=0,8 BWK_TRACE=1 python profiles/prof_fake_rank.py [ne nlev qsize reps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge

bwk = ge.build(oracle=False)
r, n = (int(x) for x in os.environ["BWK_FAKE_RANKS"].split(","))
ne, nlev, qsize, reps = (int(a) for a in (sys.argv[1:5] if len(sys.argv) >= 5 else (120, 72, 40, 4)))
nglob = 6 * ne * ne
first, count = bwk.partition(nglob, r, n)
ku = bwk.KernelUtils(count, nlev, qsize).allocate_data(host_arrays=False)
ku.initialize_data_device(nglob, first)
ku.set_topology_cubed_sphere(ne, first)
ku.set_rspheremp(__import__("numpy").ones((4, 4, count), order="F"))
for rep in range(reps):
    ku.initialize_data_device(nglob, first)
    ku.sync()
    print(f"--- rep {rep}", file=sys.stderr, flush=True)
    ku.run_biharmonic_device()
    ku.sync()
ku.finalize()
