#!/usr/bin/env python
"""bench.py -- headline benchmark of the biharmonic_wk hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--ne NE] [--nlev L] [--qsize Q]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Metric (BASELINE.json): biharmonic_wk GLL-point x level x tracer updates/s on ne120np4 L72 q40,
one update = the FULL biharmonic (Laplace -> DSS -> *rspheremp -> Laplace) at one GLL point of one
level of one tracer.  Strong scaling: the 86,400 elements are split into contiguous blocks over
the N ranks, DSS edge rows travel with ncclSend/ncclRecv.

A step = one pass of the operator over the whole (device-resident) field.  Inputs are
re-generated on the device before every step (outside the timed region: repeated application
would underflow), each step is timed with CUDA events on the library's stream, the K step
durations are summed, and the maximum over ranks is reported.  The 32 GB working set is far
larger than the 126 MB L2, so no explicit flush is needed between steps.

The same JSON line also carries
  laplace_only : the reference-pinned single Laplacian (what the reference's own
                 biharmonic_wk_scalar computes, F90:186-200) on the same data,
  roofline     : the slower of the two sweep kernels against the measured HBM copy rate,
  cpu_baseline : the CPU oracle (C port of the reference, OpenMP over elements) on a bounded
                 sample, timed on this box's host cores (N=1 only),
  e2e          : the same metric through bwk_biharmonic_wk_scalar_full with pinned HOST buffers,
                 H2D and D2H copies inside the timed region.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "biharmonic_wk GLL-pt*lev*tracer updates/s (ne120np4 L72 q40)"
UNIT = "updates/s"


def measured_hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md, 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def mark(self):
        return time.time()

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self, t0, t1):
        sm, mx, reasons = [], [], set()
        for t, line in self.rows:
            if t < t0 - 0.05 or t > t1 + 0.05:
                continue
            p = [x.strip() for x in line.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0]))
                mx.append(float(p[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def _host_threads():
    """Threads the CPU legs may use: the cores this process may run on.  torchrun exports OMP_NUM_THREADS=1 to its
    workers; the CPU arm must not inherit that (round 1's N>1 reference lines ran on one core)."""
    forced = os.environ.get("BWK_CPU_THREADS")
    n = int(forced) if forced else len(os.sched_getaffinity(0))
    os.environ["OMP_NUM_THREADS"] = str(n)  # before libbwk_oracle_omp.so (libgomp) is loaded
    return n


def _oracle():
    _host_threads()
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as orc

    orc.build()
    return orc


class CpuProblem:
    """One cubed sphere on the host for the CPU oracle (C port of F90:100-202 + DSS, OpenMP over elements)."""

    def __init__(self, orc, ne, nlev, qsize):
        self.orc, self.ne, self.nlev, self.qsize = orc, ne, nlev, qsize
        self.n = 6 * ne * ne
        self.dvv, self.elem, self.q = orc.initialize_data_parallel(self.n, nlev, qsize)
        self.dss = orc.Dss(self.n, orc.cubed_sphere_gdof(ne))
        self.rs = self.dss.rspheremp(self.elem)
        self.updates = 16.0 * nlev * qsize * self.n

    def reinit(self):
        self.orc.initialize_data_parallel(self.n, self.nlev, self.qsize, qtens=self.q)

    def step_full(self):
        t = time.perf_counter()
        self.dss.biharmonic_full(self.elem, self.rs, self.q, self.dvv, omp=True)
        return time.perf_counter() - t

    def step_laplace(self, omp=True):
        t = time.perf_counter()
        self.orc.biharmonic_wk_scalar_cpu(self.elem, self.q, self.dvv, omp=omp)
        return time.perf_counter() - t

    def label(self):
        return f"ne{self.ne}np4 L{self.nlev} q{self.qsize} ({self.n} elements)"


def _host_mem_available_gb():
    try:
        import psutil

        return psutil.virtual_memory().available / 2**30
    except Exception:
        return 0.0


def port_equals_translated_reference(orc, nlev, qsize):
    """The CPU legs time the C port (OpenMP build).  Where oracle/_ref travelled, say in the line itself whether that port's
    Laplacian equals the reference's own CPU module (F90:100-202, translated from the Fortran to C, oracle/ref_atm_py.py) bit
    for bit on the reference's default problem (F90:11-17).  None = the checker is not on this box."""
    try:
        import numpy as np
        import ref_atm_py

        if not ref_atm_py.available():
            return None
        dvv, elem, q = orc.initialize_data(16, nlev, qsize)
        a = orc.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv, omp=True)
        b = ref_atm_py.biharmonic_wk_scalar_cpu(elem, q.copy(order="F"), dvv)
        return bool(np.array_equal(a, b))
    except Exception:
        return None


def cpu_baseline(ne_small, nlev, qsize, reps, want_serial=False, ne_full=None, full_budget_s=25.0):
    """The oracle with OpenMP over elements on all host cores: a small sphere (ne_small) timed `reps` times and, when
    host memory and the time budget allow, ONE step of the headline sphere itself (ne_full)."""
    orc = _oracle()
    cores = orc.lib(omp=True).bwko_omp_max_threads()
    pr = CpuProblem(orc, ne_small, nlev, qsize)
    times, lap_times = [], []
    for _ in range(reps):
        pr.reinit()
        times.append(pr.step_full())
        pr.reinit()
        lap_times.append(pr.step_laplace())
    out = {
        "value": pr.updates / min(times), "unit": UNIT, "cores": cores, "kind": "port",
        "sample": f"{pr.label()}, full biharmonic, OpenMP over elements, best of {reps}",
        "times_s": [round(t, 4) for t in times],
        "laplace_only_value": pr.updates / min(lap_times),
        "port_laplacian_bit_identical_to_translated_reference": port_equals_translated_reference(orc, nlev, qsize),
    }
    if want_serial:
        pr.reinit()
        out["laplace_only_serial_value"] = pr.updates / pr.step_laplace(omp=False)
        out["serial_note"] = "1 core, no OpenMP: the reference's own CPU path is serial (no !$omp in atmosphere/)"
        try:  # the reference's own CPU module (F90:100-202), its Fortran translated statement by statement to C and compiled here
            import ref_atm_py

            if ref_atm_py.available():
                pr.reinit()
                t = time.perf_counter()
                ref_atm_py.biharmonic_wk_scalar_cpu(pr.elem, pr.q, pr.dvv)
                dt = time.perf_counter() - t
                out["laplace_only_reference_translated"] = {
                    "value": pr.updates / dt, "unit": UNIT, "cores": 1, "sample": pr.label(),
                    "kind": "reference (module biharmonic_wk_scalar_cpu translated to C by oracle/f90_to_c.py, gcc -O2 -ffp-contract=off; reproduces the published L2 norms)"}
        except Exception as ex:
            out["laplace_only_reference_translated"] = {"error": str(ex)[:160]}
    small_rate = out["value"]
    del pr
    if ne_full and ne_full != ne_small:
        need_gb = 128.0 * nlev * qsize * 6 * ne_full * ne_full / 2**30 * 1.15 + 2
        predicted = 16.0 * nlev * qsize * 6 * ne_full * ne_full / small_rate
        if _host_mem_available_gb() > need_gb and predicted < full_budget_s:
            pf = CpuProblem(orc, ne_full, nlev, qsize)
            t = pf.step_full()
            out["small_sample_value"] = small_rate
            out["small_sample"] = out["sample"]
            out["value"] = pf.updates / t
            out["sample"] = f"{pf.label()} = the headline workload itself, full biharmonic, OpenMP over elements, 1 step of {t:.2f} s"
            del pf
        else:
            out["full_config_skipped"] = f"ne{ne_full} needs {need_gb:.0f} GB of host memory and ~{predicted:.0f} s per step"
    return out


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path (C oracle port; the Fortran original cannot be
    compiled here) with all the host threads it can use, on rank 0 only.  Each step is one pass over the HEADLINE sphere
    (ne120 L72 q40, 32 GB on the host) when host memory allows and the K+W steps fit a few minutes; otherwise over the
    largest smaller sphere of the same L/q that does (the rate of this streaming loop does not depend on ne; said in
    `cpu_baseline.sample`)."""
    if rank != 0:
        return
    orc = _oracle()
    cores = orc.lib(omp=True).bwko_omp_max_threads()
    nsteps = args.warmup + args.steps
    # probe: one step of the small sphere -> rate -> the largest sphere whose K+W steps (plus re-initialisation) fit the budget
    probe = CpuProblem(orc, args.cpu_ne, args.nlev, args.qsize)
    probe.step_full()
    probe.reinit()
    rate = probe.updates / probe.step_full()
    del probe
    budget_s = float(os.environ.get("BWK_REF_BUDGET_S", "200"))
    mem = _host_mem_available_gb()
    ne_s, why = args.cpu_ne, ""
    for ne_c in sorted({args.ne, 90, 60, args.cpu_ne}, reverse=True):
        if ne_c > args.ne:
            continue
        upd = 16.0 * args.nlev * args.qsize * 6 * ne_c * ne_c
        need_gb = 8.0 * upd / 2**30 * 1.15 + 2
        t_step = upd / rate
        if mem > need_gb and nsteps * t_step * 1.25 < budget_s:
            ne_s = ne_c
            break
        why += f"ne{ne_c}: {need_gb:.0f} GB, {nsteps} x {t_step:.1f} s; "
    pr = CpuProblem(orc, ne_s, args.nlev, args.qsize)
    times = []
    for i in range(nsteps):
        pr.reinit()  # untimed, as the GPU arm re-generates its inputs before every step
        dt = pr.step_full()
        if i >= args.warmup:
            times.append(dt)
    tot = sum(times)
    value = pr.updates * len(times) / tot
    same = ne_s == args.ne
    sample = (f"{pr.label()}{' = the headline workload itself' if same else ' (smaller sphere, same L/q: ' + why + 'over the budget)'}, "
              f"full biharmonic, OpenMP over elements on {cores} threads, {len(times)} timed steps")
    nglob = 6 * args.ne * args.ne
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * tot / len(times), "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic (reference LCG, F90:77-91)",
        "config": workload_config(args.ne, args.nlev, args.qsize, nglob, world, args.variant),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample, "same_config": same,
                         "port_laplacian_bit_identical_to_translated_reference": port_equals_translated_reference(orc, args.nlev, args.qsize),
                         "probe_rate_small_sphere": rate, "probe_sample": f"ne{args.cpu_ne}np4 L{args.nlev} q{args.qsize}"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(ne, nlev, qsize, nglob, world, variant):
    """`config` of the JSON line, identical for both arms"""
    return {"workload": f"ne{ne}np4 L{nlev} q{qsize}, {nglob} elements, full biharmonic (Laplace -> DSS -> rspheremp -> Laplace)",
            "partition": f"contiguous element blocks over {world} rank(s)",
            "l2": "working set >> L2 (no flush needed); inputs re-generated before every step", "variant": variant}


def expected_checksum(ne, nlev, qsize):
    """bit checksum of the full biharmonic of the LCG inputs recorded from a 1-GPU run (profiles/expected_checksums.json)"""
    try:
        with open(os.path.join(ROOT, "profiles", "expected_checksums.json")) as f:
            return json.load(f).get(f"ne{ne}_L{nlev}_q{qsize}_full")
    except Exception:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--ne", type=int, default=120)
    ap.add_argument("--nlev", type=int, default=72)
    ap.add_argument("--qsize", type=int, default=40)
    ap.add_argument("--cpu-ne", type=int, default=30, help="cubed-sphere size of the CPU sample")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-hv", action="store_true", help="skip the advance_hypervis_scalar (caller, SURVEY 8(f)1) timing")
    ap.add_argument("--no-dp3d", action="store_true", help="skip the biharmonic_wk_dp3d (SURVEY 8(f)3) timing")
    ap.add_argument("--no-mini", action="store_true", help="skip the two mini-app timings (nested_loops, mmf-mpdata-tracer; SURVEY 8(f)4)")
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--pcie", action="store_true", help="also at N=1: time plain pinned H2D/D2H copies next to the e2e call")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import numpy as np
    import torch
    import torch.distributed as dist

    import __graft_entry__ as ge

    bwk = ge.build(oracle=(rank == 0))
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("gloo", rank=rank, world_size=world)  # control plane only; data plane = NCCL inside libbwk_b200

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    ne, nlev, qsize = args.ne, args.nlev, args.qsize
    nglob = 6 * ne * ne
    first, count = bwk.partition(nglob, rank, world)
    updates_global = 16.0 * nlev * qsize * nglob
    kq = nlev * qsize

    ku = bwk.KernelUtils(count, nlev, qsize).allocate_data(host_arrays=False)
    ku.set_variant(args.variant)
    if world > 1:
        idb = [bwk.KernelUtils.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(idb, src=0)
        bwk.KernelUtils.comm_init(rank, world, idb[0])
    ku.initialize_data_device(nglob, first)
    ku.set_topology_cubed_sphere(ne, first)
    ku.get_rspheremp()  # force the (host) rspheremp setup out of the timed region

    sampler = ClockSampler(local_rank).start() if rank == 0 else None

    phase_ms = {"sweep1": [], "sweep2": []}  # per timed step of the headline run (N = 1): the two sweeps, timed INSIDE the step

    def timed_steps(run, nwarm, nsteps, phases=False):
        for _ in range(nwarm):
            ku.initialize_data_device(nglob, first)
            run()
        barrier()
        t_wall0 = time.time()
        l0 = ku.launch_count()
        tot_ms = 0.0
        launches = 0
        for _ in range(nsteps):
            ku.initialize_data_device(nglob, first)  # untimed re-init (also sweeps L2)
            ku.sync()
            if world > 1:
                dist.barrier()
                ku.comm_barrier()  # device-side: the ranks' streams start the step within one NCCL latency of each other
            la = ku.launch_count()
            ku.event_record(0)
            run()
            ku.event_record(1)
            tot_ms += ku.event_elapsed_ms(0, 1)
            if phases:
                phase_ms["sweep1"].append(ku.event_elapsed_ms(0, 5))
                phase_ms["sweep2"].append(ku.event_elapsed_ms(5, 1))
            launches += ku.launch_count() - la
        barrier()
        t_wall1 = time.time()
        return max_over_ranks(tot_ms), launches, (t_wall0, t_wall1)

    # ---- headline: full biharmonic, device resident
    if world == 1:
        ku.set_phase_event(5)  # event 5 between sweep 1 and the strip sweep of every step
    tot_ms, launches, wall = timed_steps(ku.run_biharmonic_device, args.warmup, args.steps, phases=(world == 1))
    if world == 1:
        ku.set_phase_event(-1)
    ms_per_step = tot_ms / args.steps
    value = updates_global / (ms_per_step * 1e-3)

    # ---- checksum of the last timed step's result: the bit checksum does not depend on the partition (per-rank values
    #      add up mod 2^64), so the lines of N = 1, 2, 4, 8 prove "N ranks == 1 rank bit for bit" by themselves
    ck_local = (ku.checksum_bits(first),) + tuple(ku.checksum())
    if world > 1:
        ck_all = [None] * world
        dist.all_gather_object(ck_all, ck_local)
    else:
        ck_all = [ck_local]
    bits = sum(c[0] for c in ck_all) % (1 << 64)
    want = expected_checksum(ne, nlev, qsize)
    checksum = {"bits": f"{bits:016x}", "sum": sum(c[1] for c in ck_all), "sumsq": sum(c[2] for c in ck_all),
                "expected_bits": want, "match": (f"{bits:016x}" == want) if want else None,
                "note": "bits = sum over values of mix64(bits)*(2*global_index+1) mod 2^64 of the last timed step's qtens (partition-independent); "
                        "sum/sumsq are floating-point sums in rank order (last bits may depend on N); expected = 1-GPU run recorded in profiles/expected_checksums.json"}

    # ---- reference-pinned single Laplacian on the same data
    lap_ms, lap_launches, _ = timed_steps(lambda: ku.run_device(1, count), 1, args.steps)
    lap_ms /= args.steps
    peak, peak_src = measured_hbm_peak()
    bytes_sweep_local = count * (256.0 * kq + 1152.0) + 128.0  # 16 B/pt + metric
    lap = {"value": updates_global / (lap_ms * 1e-3), "unit": UNIT, "ms_per_step": lap_ms,
           "note": "one laplace_sphere_wk per update (what the reference's biharmonic_wk_scalar computes, F90:186-200)"}
    if world == 1:
        lap["roofline_frac"] = bytes_sweep_local / (lap_ms * 1e-3) / 1e9 / peak

    # ---- per-kernel timing for the roofline (single rank only: the phase hook has no halo)
    roofline = None
    kernels = {}
    if world == 1:
        for phase, name in ((1, "sweep1 (laplace -> tmp)"), (2, "sweep2 (strip DSS unpack + rspheremp + laplace)")):
            ms = []
            for _ in range(max(3, min(args.steps, 10))):
                ku.initialize_data_device(nglob, first)
                if phase == 2:
                    ku.run_biharmonic_phase(1)
                ku.sync()
                ku.event_record(2)
                ku.run_biharmonic_phase(phase)
                ku.event_record(3)
                ms.append(ku.event_elapsed_ms(2, 3))
            iso = sum(ms) / len(ms)
            inside = phase_ms["sweep1" if phase == 1 else "sweep2"]
            avg = sum(inside) / len(inside)  # the kernel's average duration over the timed region (events inside every timed step)
            kernels[name] = {"avg_ms": avg, "algorithmic_GBps": bytes_sweep_local / (avg * 1e-3) / 1e9,
                             "isolated_avg_ms": iso, "timing": "CUDA events on the library stream inside each of the timed steps; isolated = the phase launched alone after a sync"}
        dom = max(kernels, key=lambda k: kernels[k]["avg_ms"])
        ach = kernels[dom]["algorithmic_GBps"]
        roofline = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                    "kernel": dom, "peak_source": peak_src, "algorithmic_bytes_per_launch": bytes_sweep_local,
                    "step_frac_of_32B_per_update": (2 * bytes_sweep_local) / (ms_per_step * 1e-3) / 1e9 / peak,
                    "frac_of_nominal_8TBs": ach / 8000.0, "kernels": kernels}
        # DRAM bytes per launch from the committed ncu --set full capture (profiles/traffic.json, taken at a
        # smaller cubed sphere: ncu replays every kernel ~40x), scaled by the number of slabs of this launch
        # FP64 side: 64 flop per point-update per Laplacian (the reference's CPU operation count) against a
        # DFMA rate measured on this GPU; the kernels execute 40 FMA-class instructions per point (folded metric)
        try:
            fp64_peak = ku.measure_fp64_peak()
            roofline["fp64"] = {"peak_TFLOPs_measured_dfma": fp64_peak,
                                "algorithmic_TFLOPs": 2 * 64.0 * updates_global / (ms_per_step * 1e-3) / 1e12,
                                "executed_TFLOPs": 2 * 40.0 * updates_global / (ms_per_step * 1e-3) / 1e12,
                                "executed_frac_of_peak": 2 * 40.0 * updates_global / (ms_per_step * 1e-3) / 1e12 / fp64_peak}
        except Exception as ex:
            roofline["fp64"] = {"error": str(ex)[:120]}
        tr = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tr):
            try:
                tj = json.load(open(tr))
                want = "sweep2 (default form" if dom.startswith("sweep2") else "sweep_kernel<1"
                for kname, rec in tj["kernels"].items():
                    if want in kname:
                        roofline["traffic"] = rec["dram_bytes_per_slab"] * kq * count
                        roofline["traffic_note"] = (f"dram__bytes_read+write of {kname} in {tj['source']} ({tj['config']}) = "
                                                    f"{rec['dram_bytes_per_slab']:.1f} B/slab, scaled to this launch's {kq * count} slabs")
            except Exception:
                pass

    clocks = sampler.summary(*wall) if sampler else None

    # ---- the caller of the path (upstream HOMME advance_hypervis_scalar), one subcycle, device resident
    hv = None
    if not args.no_hv:
        try:
            rng = np.random.default_rng(1 + rank)
            dp = np.asfortranarray(0.5 + rng.random((4, 4, nlev, count)))
            ku.upload_hypervis_fields(dp, np.asfortranarray(dp[::-1]))
            del dp
            hv_ms, hv_launches, _ = timed_steps(lambda: ku.run_advance_hypervis_device(2.0, 1.0e24, 1, True), 1, max(3, min(args.steps, 5)))
            hv_ms /= max(3, min(args.steps, 5))
            hv = {"value": updates_global / (hv_ms * 1e-3), "unit": UNIT, "ms_per_step": hv_ms, "bytes_per_update": 56,
                  "GBps_of_its_traffic": 56.0 * updates_global / world / (hv_ms * 1e-3) / 1e9,
                  "note": "Qtens=dpdiss*Qdp/dp -> biharmonic -> Qdp*spheremp-dt*nu_q*Qtens -> limiter2d_zero -> DSS -> *rspheremp (3 launches: sweep 1, strip sweep with the update fused in, DSS-only sweep)"}
        except Exception as ex:
            hv = {"value": None, "error": str(ex)[:200]}

    # ---- e2e: host buffers through the drop-in call (H2D + D2H inside the timed region)
    e2e = None
    if not args.no_e2e:
        es = 16 * kq
        try:
            hq = ku._pinned_array((4, 4, nlev, qsize, count))
            helem = np.zeros(count, dtype=bwk.ELEMENT_DTYPE)
            hderiv = np.zeros((), dtype=bwk.DERIVATIVE_DTYPE)
            # host inputs = the device-generated LCG data (reference generator), copied back once
            import ctypes as C

            def refill():
                ku.initialize_data_device(nglob, first)
                bwk.binding._ck(bwk.lib().bwk_download(hq.ctypes.data_as(C.c_void_p)))

            # element metric on the host from the closed-form LCG (F90:52-56)
            state, tab = 11, np.empty(131072)
            for t in range(131072):
                state = (1301 * state + 97) % 131072
                tab[t] = state / 131072.0
            d = 16 + 144 * (first + np.arange(count, dtype=np.int64))[:, None] + np.arange(144, dtype=np.int64)[None, :]
            helem.view(np.float64).reshape(count, 144)[...] = tab[d % 131072]
            hderiv["Dvv"] = tab[:16].reshape(4, 4)
            if world > 1:
                # with host-supplied element records a rank cannot see its neighbours' spheremp:
                # hand the library its rspheremp explicitly (HOMME keeps elem%rspheremp per element)
                ku.set_rspheremp(ku.get_rspheremp())
            tot = 0.0
            nst = max(1, args.e2e_steps)
            for i in range(1 + nst):
                refill()
                barrier()
                t0 = time.perf_counter()
                ku.biharmonic_wk_scalar_full(helem, hq, hderiv)
                torch.cuda.synchronize()
                dt = time.perf_counter() - t0
                if i >= 1:
                    tot += max_over_ranks(dt)
            bytes_q = 8.0 * es * count
            pcie = None
            if world > 1 or args.pcie:
                # what the host side of this box gives every rank when all ranks copy at once: the
                # e2e call cannot go faster than min(h2d, d2h) together
                nb = min(int(bytes_q), 4 << 30)
                hbuf = torch.from_numpy(hq.reshape(-1, order="F")[: nb // 8])
                dbuf = torch.empty(nb // 8, dtype=torch.float64, device="cuda")
                dbuf2 = torch.empty(nb // 8, dtype=torch.float64, device="cuda")
                s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

                def timed_copy(up, down):
                    barrier()
                    t0 = time.perf_counter()
                    if up:
                        with torch.cuda.stream(s1):
                            dbuf.copy_(hbuf, non_blocking=True)
                    if down:
                        with torch.cuda.stream(s2):
                            hbuf2.copy_(dbuf2, non_blocking=True)
                    torch.cuda.synchronize()
                    return max_over_ranks(time.perf_counter() - t0)

                hbuf2 = torch.from_numpy(hq.reshape(-1, order="F")[nb // 8: 2 * (nb // 8)]) if 2 * nb <= bytes_q else torch.empty(nb // 8, dtype=torch.float64).pin_memory()
                timed_copy(True, True)
                pcie = {"bytes_per_copy": nb, "ranks_concurrent": world,
                        "h2d_alone_GBps_per_rank": nb / timed_copy(True, False) / 1e9,
                        "d2h_alone_GBps_per_rank": nb / timed_copy(False, True) / 1e9,
                        "both_GBps_per_rank_per_direction": nb / timed_copy(True, True) / 1e9,
                        "note": "all ranks copy at the same time from/to pinned host memory (cudaMemcpyAsync through torch); slowest rank"}
                del dbuf, dbuf2
            e2e = {"value": updates_global * nst / tot, "unit": UNIT, "h2d_bytes_per_step": bytes_q + 1152.0 * count + 128,
                   "d2h_bytes_per_step": bytes_q, "ms_per_step": 1e3 * tot / nst, "steps": nst,
                   "api": "bwk_biharmonic_wk_scalar_full(elem, qtens, deriv) with pinned host arrays, host wall clock",
                   "GBps_per_rank_per_direction": bytes_q / (tot / nst) / 1e9, "pcie": pcie}
        except Exception as ex:  # e.g. not enough pinned host memory
            e2e = {"value": None, "unit": UNIT, "error": str(ex)[:200]}

    if sampler:
        sampler.stop()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline(args.cpu_ne, nlev, qsize, 3, want_serial=True, ne_full=ne)

    # ---- biharmonic_wk_dp3d: T, v, dp3d of every level, its own context (qsize = 4), device resident
    dp3d = None
    if world == 1 and not args.no_dp3d:
        ku.finalize()
        try:
            kd = bwk.KernelUtils(count, nlev, 4).allocate_data(host_arrays=False)
            kd.initialize_data_device(nglob, first)
            kd.set_topology_cubed_sphere(ne, first)
            kd.get_rspheremp()
            kd.upload_vec_sphere2cart(np.random.default_rng(1).random((count, 96)))
            nd = max(3, min(args.steps, 5))
            tot = 0.0
            for i in range(1 + nd):
                kd.initialize_data_device(nglob, first)
                kd.sync()
                l0 = kd.launch_count()
                kd.event_record(0)
                kd.run_biharmonic_dp3d_device(True, True)
                kd.event_record(1)
                kd.sync()
                if i >= 1:
                    tot += kd.event_elapsed_ms(0, 1)
                    d_launches = kd.launch_count() - l0
            d_ms = tot / nd
            pts = 16.0 * nlev * 4 * count
            dp3d = {"value": pts / (d_ms * 1e-3), "unit": "field-point updates/s (4 fields per level: T, v(1), v(2), dp3d)", "ms_per_step": d_ms,
                    "bytes_per_update": 48, "GBps_of_its_traffic": 48.0 * pts / (d_ms * 1e-3) / 1e9, "launches_per_step": d_launches,
                    "note": "laplace_sphere_wk(T, dp3d) + Cartesian vlaplace_sphere_wk(v) -> DSS -> *rspheremp -> the same operators again "
                            "(3 launches: element-local sweep, DSS strip sweep, element-local sweep in place; a fused form would need 32 B/update)"}
            kd.finalize()
        except Exception as ex:
            dp3d = {"value": None, "error": str(ex)[:200]}

    # ---- the reference repo's two other mini-apps, each in its own process (own library, own
    #      CUDA context; a failure there cannot touch this line's numbers), bounded to two minutes each
    mini = None
    if rank == 0 and world == 1 and not args.no_mini:
        mini = {}
        for name, script in (("nested_loops", "prof_nested.py"), ("mmf_mpdata_tracer", "prof_mpdata.py")):
            try:
                r = subprocess.run([sys.executable, os.path.join(ROOT, "profiles", script)], capture_output=True, text=True, timeout=120)
                last = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
                mini[name] = json.loads(last[-1]) if r.returncode == 0 and last else {"error": (r.stderr or r.stdout)[-300:]}
            except Exception as ex:
                mini[name] = {"error": str(ex)[:200]}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic (reference LCG F90:77-91, generated on the device; cubed-sphere connectivity SURVEY App. D)",
            "config": workload_config(ne, nlev, qsize, nglob, world, args.variant), "checksum": checksum,
            "laplace_only": lap, "advance_hypervis_scalar": hv, "biharmonic_wk_dp3d": dp3d, "mini_apps": mini, "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    ku.finalize()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
