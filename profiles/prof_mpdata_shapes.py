"""GPU-box script (run under gpurun): launch-shape A/B of the MPDATA step (libmpdata_b200.so) at one slice count.

    python profiles/prof_mpdata_shapes.py [--slices 9472] [--shapes 2x512,2x320,3x256] [--iters 20] > gpurun_out/x.json

A shape is <level chunks per slice>x<threads per CTA> (0 = the library's choice) for the barrier-phased kernel, or m<warps per
CTA> for the marching kernel (mpdata_march.cuh; m0 = the library's choice), d<warps per CTA> for the marching kernel with two
slices per warp.  Prints one JSON object: per shape the mean
kernel time (CUDA events on the library stream), the launch the library made of it (grid, CTAs per SM follow from it) and
whether the result of one step equals the oracle's bit for bit.  The oracle is used here only as the checker.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--slices", type=int, default=9472)
    ap.add_argument("--shapes", default="0x0,2x512,2x320,2x256,2x384,3x256,3x320,4x192,4x256,1x1024,1x512")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--no-check", action="store_true")
    args = ap.parse_args()
    import __graft_entry__ as ge

    ge.build()
    from codesign_kernels_b200 import mpdata
    import mpdata_oracle_py as mpo

    nx, nz, ns = mpdata.NX, mpdata.NZ, args.slices
    P = mpdata.TestAdvect(ns, nx, nz).init(seed=100)
    ref = None
    if not args.no_check:
        f, fl = P.f.copy(order="F"), P.flux.copy(order="F")
        mpo.advect_scalar2D_cpu(ns, nx, nz, f, P.u, P.w, P.rho, P.rhow, P.adz, fl, omp=True)
        ref = (f, fl)
    b = mpdata.algorithmic_bytes(ns, nx, nz)
    peak = 6553.6
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    out = {"nslices": ns, "nx": nx, "nz": nz, "algorithmic_bytes": b, "hbm_peak_GBs": peak, "env": {k: v for k, v in os.environ.items() if k.startswith("MPDATA_B200")}, "shapes": {}}
    for s in args.shapes.split(","):
        try:
            if s[0] in "md":
                ms = mpdata.time_steps(P, iters=args.iters, block=0, chunks=0, form=2 if s[0] == "m" else 3, warps_per_cta=int(s[1:]))
            else:
                c_, b_ = (int(x) for x in s.split("x"))
                ms = mpdata.time_steps(P, iters=args.iters, block=b_, chunks=c_, form=1 if (c_ or b_) else 0)
        except Exception as e:  # a shape the library refuses (shared memory)
            out["shapes"][s] = {"error": str(e)}
            continue
        rec = {"ms_per_step": ms, "frac_of_hbm": b / (ms * 1e-3) / 1e9 / peak, "launch": dict(zip(("grid", "block", "chunks", "smem_bytes"), mpdata.launch_shape()))}
        if ref is not None:
            fo, flo = np.empty_like(P.f), P.flux.copy(order="F")
            assert mpdata.lib().mpdata_download(fo.ctypes.data, flo.ctypes.data) == 0
            # time_steps restarts every step from the uploaded field: what is on the device is ONE step
            rec["bitwise_equal_to_oracle"] = bool(np.array_equal(fo, ref[0]) and np.array_equal(flo, ref[1]))
        out["shapes"][s] = rec
    print(json.dumps(out))


if __name__ == "__main__":
    main()
