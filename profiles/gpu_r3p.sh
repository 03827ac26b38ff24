#!/bin/bash
# r3p: the MPDATA GPU tests with the automatic form choice (mpdata_march_wins) in force, then the choice at three slice counts
mkdir -p gpurun_out
timeout 24 python -m pytest tests/test_zzz_mpdata.py -m gpu -x -q > gpurun_out/r3p_pytest_mpdata.txt 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r3p_pytest_mpdata.txt
for ns in 1600 2600; do
  timeout 5 python profiles/prof_mpdata_shapes.py --slices $ns --shapes 2x512,m6,0x0 > gpurun_out/r3p_mpdata_shapes_$ns.json 2>> gpurun_out/r3p_mpdata_shapes.err; echo "shapes $ns rc $?"
done
python - <<'PY'
import json
for ns in (1600, 2600):
    try:
        d=json.load(open(f"gpurun_out/r3p_mpdata_shapes_{ns}.json"))
        for k,v in d["shapes"].items(): print(ns, k, v.get("ms_per_step"), v.get("launch"), v.get("bitwise_equal_to_oracle"), v.get("error"))
    except Exception as e: print(ns, "no result", e)
PY
