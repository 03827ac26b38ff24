"""Turns an .ncu-rep (ncu --set full) into the small per-kernel CSV committed under profiles/ and,
optionally, profiles/traffic.json (DRAM bytes per launch and per slab, read by bench.py).

usage: python profiles/summarize_ncu.py gpurun_out/X.ncu-rep profiles/X.csv [--traffic NE NLEV QSIZE]
Runs in the build container (ncu -i needs no GPU)."""
import csv
import io
import json
import os
import subprocess
import sys

KEEP = (
    "Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "gpu__time_duration.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__cycles_active.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sectors_srcunit_tex_op_read.sum", "l1tex__m_xbar2l1tex_read_bytes.sum", "l1tex__m_l1tex2xbar_write_bytes.sum",
    "l1tex__m_xbar2l1tex_read_bytes_mem_global_op_tma_ld.sum", "l1tex__m_l1tex2xbar_write_bytes_mem_global_op_tma_st.sum",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__cycles_elapsed.avg",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smsp__sass_inst_executed_op_tma_ld.sum", "smsp__sass_inst_executed_op_tma_st.sum",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
)


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["metric", "unit"] + [f"launch{i}" for i in range(len(data))])
        for k in KEEP:
            if k in hdr:
                i = hdr.index(k)
                w.writerow([k, units[i]] + [r[i] for r in data])
    print(open(out).read())
    if "--traffic" in sys.argv:
        a = sys.argv.index("--traffic")
        ne, nlev, qsize = (int(x) for x in sys.argv[a + 1 : a + 4])
        slabs = 6 * ne * ne * nlev * qsize
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
        ir, iw, ik = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("Kernel Name")
        tr = {"source": os.path.basename(rep), "config": f"ne{ne} L{nlev} q{qsize}", "slabs_per_launch": slabs, "kernels": {}}
        for r in data:
            b = float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]]
            tr["kernels"][r[ik]] = {"dram_bytes_per_launch": b, "dram_bytes_per_slab": b / slabs, "algorithmic_bytes_per_slab": 256.0}
        json.dump(tr, open(os.path.join(os.path.dirname(out), "traffic.json"), "w"), indent=1)
        print(json.dumps(tr, indent=1))


if __name__ == "__main__":
    main()
