"""CPU: the JSON line of `bench.py --impl reference` (the CPU arm the driver runs beside ours) at a tiny size -- the keys the
contract names, and that under a multi-rank launch only rank 0 prints.  The GPU arm's line cannot be produced without a GPU;
its keys are the same plus `roofline`, `gpu_launches`, `clocks` (profiles/r3i_bench.json is a kept example, checked here too)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data",
             "config", "cpu_baseline", "e2e"}


def _run(extra_env=None, args=()):
    env = dict(os.environ, BWK_REF_BUDGET_S="60")
    env.update(extra_env or {})
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--ne", "4", "--cpu-ne", "2", *args],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stderr[-400:]
    return [ln for ln in r.stdout.splitlines() if ln.startswith("{")]


def test_reference_arm_line_has_the_contract_keys():
    lines = _run()
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and BASE_KEYS <= set(d)
    assert d["unit"] == "updates/s" and d["higher_is_better"] is True and d["dtype"] == "f64" and d["vs_baseline"] is None
    assert "workload" in d["config"] and "ne4np4" in d["config"]["workload"] and "model" not in d["config"]
    cb = d["cpu_baseline"]
    assert {"value", "unit", "cores", "kind", "sample"} <= set(cb) and cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"]
    assert cb["port_laplacian_bit_identical_to_translated_reference"] in (True, None)  # None: oracle/_ref is not on this box
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["value"] > 0 and abs(d["ms_per_step"] * 1e-3 * d["value"] - 16 * 72 * 40 * 96) < 1e-6 * 16 * 72 * 40 * 96


def test_reference_arm_other_ranks_stay_silent():
    assert _run({"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2"}, ("--gpus", "2")) == []


def test_kept_gpu_line_has_the_contract_keys():
    d = json.load(open(os.path.join(ROOT, "profiles", "r3i_bench.json")))
    assert BASE_KEYS | {"roofline", "gpu_launches", "clocks"} <= set(d) and d["gpu_launches"] > 0
    rf = d["roofline"]
    assert {"bound", "achieved", "peak", "unit", "frac", "traffic"} <= set(rf) and rf["bound"] == "hbm" and abs(rf["frac"] - rf["achieved"] / rf["peak"]) < 1e-9
    assert {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"} <= set(d["e2e"]) and d["e2e"]["h2d_bytes_per_step"] > 0
    assert d["checksum"]["match"] is True
