"""PCIe probes + e2e timing of bwk_biharmonic_wk_scalar_full at ne120 (or given ne)."""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as ge
bwk = ge.build(oracle=False)
ne = int(sys.argv[1]) if len(sys.argv) > 1 else 120
nlev, qsize = 72, 40
n = 6 * ne * ne
nbytes = n * 16 * nlev * qsize * 8
if "pcie" in sys.argv:
    h1 = torch.empty(nbytes // 4, dtype=torch.uint8).pin_memory()
    h2 = torch.empty(nbytes // 4, dtype=torch.uint8).pin_memory()
    d1 = torch.empty(nbytes // 4, dtype=torch.uint8, device="cuda")
    d2 = torch.empty(nbytes // 4, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for name in ("h2d", "d2h", "both"):
        torch.cuda.synchronize(); t = time.perf_counter()
        if name in ("h2d", "both"):
            with torch.cuda.stream(s1): d1.copy_(h1, non_blocking=True)
        if name in ("d2h", "both"):
            with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
        torch.cuda.synchronize(); dt = time.perf_counter() - t
        print(f"{name}: {nbytes/4/dt/1e9:.1f} GB/s per direction ({dt*1e3:.1f} ms for {nbytes/4/1e9:.1f} GB)")
    del h1, h2, d1, d2
ku = bwk.KernelUtils(n, nlev, qsize).allocate_data(host_arrays=False)
ku.initialize_data_device()
ku.set_topology_cubed_sphere(ne)
ku.get_rspheremp()
import ctypes as C
hq = ku._pinned_array((4, 4, nlev, qsize, n))
helem = np.zeros(n, dtype=bwk.ELEMENT_DTYPE)
hderiv = np.zeros((), dtype=bwk.DERIVATIVE_DTYPE)
state, tab = 11, np.empty(131072)
for t in range(131072):
    state = (1301 * state + 97) % 131072
    tab[t] = state / 131072.0
d = 16 + 144 * np.arange(n, dtype=np.int64)[:, None] + np.arange(144, dtype=np.int64)[None, :]
helem.view(np.float64).reshape(n, 144)[...] = tab[d % 131072]
hderiv["Dvv"] = tab[:16].reshape(4, 4)
for i in range(3):
    ku.initialize_data_device()
    bwk.binding._ck(bwk.lib().bwk_download(hq.ctypes.data_as(C.c_void_p)))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ku.biharmonic_wk_scalar_full(helem, hq, hderiv)
    torch.cuda.synchronize()
    print(f"e2e full rep {i}: {(time.perf_counter()-t0)*1e3:.1f} ms")
for i in range(2):
    ku.initialize_data_device()
    bwk.binding._ck(bwk.lib().bwk_download(hq.ctypes.data_as(C.c_void_p)))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ku.biharmonic_wk_scalar(helem, hq, None, hderiv, 1, n)
    torch.cuda.synchronize()
    print(f"e2e single laplace rep {i}: {(time.perf_counter()-t0)*1e3:.1f} ms")
ku.finalize()
