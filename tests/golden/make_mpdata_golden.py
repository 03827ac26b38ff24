"""Generates tests/golden/mpdata_results.json.  Run in the build container (needs /root/reference):

    python tests/golden/make_mpdata_golden.py

(1) What the reference holds for the mmf-mpdata-tracer kernel: the lines its driver printed on
    Titan under mmf-mpdata-tracer/results/ -- timings and its acceptance quantities, the relative
    L1 errors of f and flux of each OpenACC form against advect_scalar2D_cpu
    (advect_scalar2D_pushncols_openacc.F90:679-683).  No arrays are stored there.
(2) A regression pin of the C oracle (NOT reference output): SHA-256 of f and flux after one step
    of oracle/mpdata_oracle.c on inputs from the integer LCG below (no numpy RNG, no libm), at the
    reference sizes.  It changes if anybody edits the oracle's arithmetic.
(3) The same digest from the reference's own advect_scalar2D_cpu, mechanically translated to C from
    the source under /root/reference (oracle/f90_to_c.py, `make -C oracle ref`): the two are equal,
    which makes (2) a pin on the reference's arithmetic.
"""
import glob
import hashlib
import json
import os
import re
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = "/root/reference/mmf-mpdata-tracer/results"
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def lcg_fields(nslices=48, nx=32, nz=58):
    """Deterministic inputs in the reference's draw order (adz, f, u, w, rho, rhow, flux; A:654-660):
    x <- (1103515245 x + 12345) mod 2^31, value = (x >> 8) / 2^23 in [0,1); u and w shifted to [-0.5,0.5)."""
    from codesign_kernels_b200 import mpdata

    sh = mpdata.shapes(nslices, nx, nz)
    order = ("adz", "f", "u", "w", "rho", "rhow", "flux")
    n = sum(int(np.prod(sh[k])) for k in order)
    x = np.empty(n, dtype=np.int64)
    s = 100  # the reference's seed value (A:650), used here as the LCG state
    for j in range(n):
        s = (1103515245 * s + 12345) % (1 << 31)
        x[j] = s
    vals = (x >> 8).astype(np.float64) / float(1 << 23)
    out, o = {}, 0
    for k in order:
        m = int(np.prod(sh[k]))
        a = vals[o:o + m].reshape(sh[k], order="F")
        if k in ("u", "w"):
            a = a - 0.5
        if k in ("adz", "rho", "rhow"):
            a = 0.5 + a  # keep the divisors away from 0
        out[k] = np.asfortranarray(a)
        o += m
    return out


def oracle_digest():
    import mpdata_oracle_py as mpo

    a = lcg_fields()
    f, fl = a["f"].copy(order="F"), a["flux"].copy(order="F")
    mpo.advect_scalar2D_cpu(48, 32, 58, f, a["u"], a["w"], a["rho"], a["rhow"], a["adz"], fl)
    return {"f_sha256": hashlib.sha256(f.tobytes(order="F")).hexdigest(), "flux_sha256": hashlib.sha256(fl.tobytes(order="F")).hexdigest(),
            "f_sum": float(f.sum()), "flux_sum_levels_1_nzm": float(fl[:, :-1].sum())}


def translated_reference_digest():
    """The same digest from the REFERENCE'S OWN advect_scalar2D_cpu (A:477-642), translated mechanically by
    oracle/f90_to_c.py from the source under /root/reference and compiled here (oracle/ref_mpdata_py.py)."""
    import ref_mpdata_py as ref

    a = lcg_fields()
    f, fl = a["f"].copy(order="F"), a["flux"].copy(order="F")
    ref.advect_scalar2D_cpu(48, 32, 58, f, a["u"], a["w"], a["rho"], a["rhow"], a["adz"], fl)
    return {"f_sha256": hashlib.sha256(f.tobytes(order="F")).hexdigest(), "flux_sha256": hashlib.sha256(fl.tobytes(order="F")).hexdigest(),
            "f_sum": float(f.sum()), "flux_sum_levels_1_nzm": float(fl[:, :-1].sum())}


def main():
    files = {}
    for path in sorted(glob.glob(os.path.join(REF, "advect.*"))):
        name = os.path.basename(path)
        if name.endswith("nvprof"):
            continue
        rows = {"cpu_timing_s": [], "openacc_timing_s": [], "rel_l1_f": [], "rel_l1_flux": []}
        for line in open(path, errors="replace"):
            m = re.match(r"\s*(CPU Timing|OpenACC-\d Timing|Relative L1 Error - f\s*|Relative L1 Error - flux\s*):\s*([-+0-9.Ee]+)", line)
            if not m:
                continue
            key, val = m.group(1).strip(), float(m.group(2))
            rows["cpu_timing_s" if key == "CPU Timing" else "openacc_timing_s" if key.startswith("OpenACC") else
                 "rel_l1_f" if key.endswith("- f") else "rel_l1_flux"].append(val)
        if any(rows.values()):
            files[name] = rows
    out = {"source": "mmf-mpdata-tracer/results/advect.* (driver output on Titan; sizes nslices=48 nz=58 nx=32, A:6-8)",
           "reference_results": files,
           "acceptance": {"note": "relative L1 error against advect_scalar2D_cpu as printed by the reference's own builds: exactly 0 only for "
                                  "gnu.7.1.0; other CPU builds of the OpenACC-annotated forms differ by up to 1e-4 (intel), 1e-6 (pgi), 5e-8 (gnu.4.9.3)",
                          "worst_cpu_build_rel_l1_f": max(v for n, r in files.items() if n != "advect.pgiacc.17.7" for v in r["rel_l1_f"] + [0.0]),
                          "worst_cpu_build_rel_l1_flux": max(v for n, r in files.items() if n != "advect.pgiacc.17.7" for v in r["rel_l1_flux"] + [0.0]),
                          "gpu_build": "advect.pgiacc.17.7",
                          "gpu_build_rel_l1_f": max(files["advect.pgiacc.17.7"]["rel_l1_f"]),
                          "gpu_build_rel_l1_flux": max(files["advect.pgiacc.17.7"]["rel_l1_flux"])},
           "oracle_regression_pin": dict(oracle_digest(), note="NOT reference output: builder's C restatement on LCG inputs"),
           "translated_reference_pin": dict(translated_reference_digest(),
                                            note="output of the reference's own advect_scalar2D_cpu (A:477-642) on the same LCG inputs, the "
                                                 "Fortran translated statement by statement by oracle/f90_to_c.py and compiled with gcc -O2 "
                                                 "-ffp-contract=off (no Fortran compiler in the image); equal to the oracle's pin")}
    with open(os.path.join(ROOT, "tests", "golden", "mpdata_results.json"), "w") as fh:
        json.dump(out, fh, indent=1)
    print(json.dumps(out["acceptance"]), json.dumps(out["oracle_regression_pin"]))


if __name__ == "__main__":
    main()
