"""Timing / profiling driver of biharmonic_wk_dp3d (device resident), default ne120 L72 (4 fields per level).
usage: python profiles/prof_dp3d.py [ne nlev]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge

bwk = ge.build(oracle=False)
ne, nlev = (int(a) for a in (sys.argv[1:3] if len(sys.argv) >= 3 else (120, 72)))
n = 6 * ne * ne
ku = bwk.KernelUtils(n, nlev, 4).allocate_data(host_arrays=False)
ku.initialize_data_device()
ku.set_topology_cubed_sphere(ne)
ku.get_rspheremp()
rng = np.random.default_rng(1)
ku.upload_vec_sphere2cart(rng.random((n, 96)))
for rep in range(4):
    ku.initialize_data_device()
    ku.sync()
    l0 = ku.launch_count()
    ku.event_record(0); ku.run_biharmonic_dp3d_device(True, True); ku.event_record(1)
    ku.sync()
    ms = ku.event_elapsed_ms(0, 1)
    pts = 16.0 * nlev * 4 * n
    # bytes this implementation must move per point and field: sweep (8 r + 8 w) + DSS sweep 16 + sweep in place 16 = 48;
    # a fused form (DSS + second operator in one strip sweep, as the scalar path has) would need 32
    print(f"rep {rep} biharmonic_wk_dp3d: {ms:.3f} ms  {pts / ms / 1e6:.1f} G field-point updates/s  {48 * pts / ms / 1e6:.1f} GB/s of its 48 B/update "
          f"({32 * pts / ms / 1e6:.1f} GB/s of the 32 B/update a fused form needs)  launches {ku.launch_count() - l0}")
ku.finalize()
