"""Profiling driver: one warm-up and one measured launch of each sweep kernel (laplace, sweep1,
sweep2) on a cubed sphere small enough for ncu's replay (default ne30 L72 q40, 2 GB).
usage: python profiles/prof_run.py [ne nlev qsize]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge

bwk = ge.build(oracle=False)
ne, nlev, qsize = (int(a) for a in (sys.argv[1:4] if len(sys.argv) >= 4 else (30, 72, 40)))
n = 6 * ne * ne
ku = bwk.KernelUtils(n, nlev, qsize).allocate_data(host_arrays=False)
ku.initialize_data_device()
ku.set_topology_cubed_sphere(ne)
ku.get_rspheremp()
for rep in range(5):
    ku.initialize_data_device()
    ku.sync()
    ku.event_record(0); ku.run_device(1, n); ku.event_record(1)
    ku.initialize_data_device()
    ku.event_record(2); ku.run_biharmonic_phase(1); ku.event_record(3)
    ku.run_biharmonic_phase(2); ku.event_record(4)
    ku.sync()
    b = n * (256.0 * nlev * qsize + 1152) + 128
    for name, a, c in (("laplace", 0, 1), ("sweep1", 2, 3), ("sweep2", 3, 4)):
        ms = ku.event_elapsed_ms(a, c)
        print(f"rep {rep} {name}: {ms:.4f} ms  {b / ms / 1e6:.1f} GB/s algorithmic")
ku.finalize()
