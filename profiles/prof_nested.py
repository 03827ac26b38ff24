"""Timing / profiling driver of the nested_loops flux kernel (libcke_b200.so) at the namelist sizes
(nested_loops/nested.nml:2-6; nvldim 104 = the GPU target's F90_PACK_SIZE=8).  No torch import.

  python profiles/prof_nested.py            -> one JSON line: warm (L2-resident, back-to-back) and cold
                                               (L2 flushed before every iteration) times per iteration,
                                               algorithmic GB/s, CPU baseline (oracle port, OpenMP over edges)
  python profiles/prof_nested.py --quick    -> a few iterations only (what runs under ncu)
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import numpy as np  # noqa: E402

from codesign_kernels_b200 import nested  # noqa: E402


def sweep():
    """A/B of the launch shape on one box: CKE_B200_UNROLL x CKE_B200_EDGES_PER_CTA x CKE_B200_GRID, every
    configuration checked bit for bit against the oracle's cke_impl1 before it is timed."""
    import cke_oracle_py as cko

    prob = nested.NestedProblem(packn=8, seed=11)
    want, _ = cko.impl1(prob)
    rows = []
    for unroll, epc, grid in [(1, 8, "persist"), (1, 8, "full"), (5, 8, "full"), (5, 0, "full"), (5, 0, "persist"),
                              (10, 0, "full"), (5, 4, "full"), (10, 8, "full")]:
        os.environ["CKE_B200_UNROLL"] = str(unroll)
        os.environ["CKE_B200_GRID"] = grid
        if epc:
            os.environ["CKE_B200_EDGES_PER_CTA"] = str(epc)
        else:
            os.environ.pop("CKE_B200_EDGES_PER_CTA", None)
        nl = nested.NestedLoops()
        nl.kokkos_init()
        nl.cke_init(prob, nIters=2)
        try:
            nl.cke_impl1_run()
            same = bool(np.array_equal(nl.cke_get_results(np.zeros((prob.nvldim, prob.nEdges), order="F")), want))
            row = {"unroll": unroll, "edges_per_cta": epc or "auto", "grid": grid, "bit_identical": same,
                   "impl1_warm_us": 1e3 * nl.time_ms(1, 300), "impl1_cold_us": 1e3 * nl.time_cold_ms(1, 21),
                   "impl2_warm_us": 1e3 * nl.time_ms(2, 300), "impl2_cold_us": 1e3 * nl.time_cold_ms(2, 21)}
        finally:
            nl.cke_cleanup()
        rows.append(row)
        print(json.dumps(row), flush=True)
    return rows


def main():
    if "--sweep" in sys.argv:
        sweep()
        return
    quick = "--quick" in sys.argv
    prob = nested.NestedProblem(packn=8, seed=11)
    nl = nested.NestedLoops()
    nl.kokkos_init()
    nl.cke_init(prob, nIters=1 if quick else 3)
    try:
        nl.cke_impl1_run()
        nl.cke_impl2_run()
        if quick:
            return
        out = {"workload": f"nested_loops nEdges={prob.nEdges} nCells={prob.nCells} nVertLevels={prob.nVertLevels} "
                           f"nvldim={prob.nvldim} nAdv={prob.nAdv}",
               "points": prob.points(), "algorithmic_bytes": prob.algorithmic_bytes()}
        for impl in (1, 2):
            warm = nl.time_ms(impl, 500)
            cold = nl.time_cold_ms(impl, 41)
            out[f"impl{impl}"] = {
                "warm_ms_per_iter": warm, "cold_ms_per_iter": cold,
                "warm_points_per_s": prob.points() / (warm * 1e-3),
                "cold_points_per_s": prob.points() / (cold * 1e-3),
                "warm_alg_GBs": prob.algorithmic_bytes() / (warm * 1e-3) / 1e9,
                "cold_alg_GBs": prob.algorithmic_bytes() / (cold * 1e-3) / 1e9,
            }
        peaks = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks):
            hbm = json.load(open(peaks)).get("hbm_gbs")
            if hbm:
                out["hbm_peak_GBs"] = hbm
                out["impl1"]["cold_frac_of_hbm"] = out["impl1"]["cold_alg_GBs"] / hbm
        import cke_oracle_py as cko

        sec, ref = cko.time_original_cpu(prob, 10)
        got = nl.cke_get_results(np.zeros((prob.nvldim, prob.nEdges), order="F"))
        bad, worst = cko.check(got, ref, prob, nested.ERR_TOL)
        out["cpu_baseline"] = {"value": prob.points() / sec, "unit": "points/s", "cores": os.cpu_count(), "kind": "port",
                               "sample": "10 iterations of the original CPU form (nested.F90:119-157), OpenMP over edges",
                               "ms_per_iter": sec * 1e3}
        out["errTol_failures_vs_cpu"] = bad
        out["worst_rel_err"] = worst
        print(json.dumps(out))
    finally:
        nl.cke_cleanup()
        nl.kokkos_finalize()


if __name__ == "__main__":
    main()
