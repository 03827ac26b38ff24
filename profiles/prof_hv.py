"""Timing / profiling driver of advance_hypervis_scalar (device resident), default ne120 L72 q40, one subcycle.
usage: python profiles/prof_hv.py [ne nlev qsize]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge

bwk = ge.build(oracle=False)
ne, nlev, qsize = (int(a) for a in (sys.argv[1:4] if len(sys.argv) >= 4 else (120, 72, 40)))
n = 6 * ne * ne
ku = bwk.KernelUtils(n, nlev, qsize).allocate_data(host_arrays=False)
ku.initialize_data_device()
ku.set_topology_cubed_sphere(ne)
ku.get_rspheremp()
rng = np.random.default_rng(1)
dp = np.asfortranarray(0.5 + rng.random((4, 4, nlev, n)))
ku.upload_hypervis_fields(dp, np.asfortranarray(dp[::-1].copy()))
for rep in range(3):
    ku.initialize_data_device()
    ku.sync()
    l0 = ku.launch_count()
    ku.event_record(0); ku.run_advance_hypervis_device(2.0, 1.0e24, 1, True); ku.event_record(1)
    ku.sync()
    ms = ku.event_elapsed_ms(0, 1)
    pts = 16.0 * nlev * qsize * n
    # bytes this implementation must move per point: sweep1 16 + sweep2 with the fused update 24 + DSS-only 16 = 56
    print(f"rep {rep} advance_hypervis_scalar: {ms:.3f} ms  {pts / ms / 1e6:.1f} G updates/s  {56 * pts / ms / 1e6:.1f} GB/s of its 56 B/update  launches {ku.launch_count() - l0}")
ku.finalize()
