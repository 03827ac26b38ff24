#!/bin/bash
# r3n: MPDATA marching forms against the barrier-phased default, timing + bitwise check at 9,472 and 1,184 slices (the r3k / r3l
# outputs were lost with a container)
mkdir -p gpurun_out
timeout 45 python profiles/prof_mpdata_shapes.py --slices 9472 --shapes 0x0,m4,m6,m12,d3,d6 > gpurun_out/r3n_mpdata_shapes_9472.json 2> gpurun_out/r3n_mpdata_shapes.err; echo "shapes 9472 rc $?"
timeout 25 python profiles/prof_mpdata_shapes.py --slices 1184 --shapes 0x0,m6,d6 > gpurun_out/r3n_mpdata_shapes_1184.json 2>> gpurun_out/r3n_mpdata_shapes.err; echo "shapes 1184 rc $?"
python - <<'PY'
import json
for ns in (9472, 1184):
    try:
        d=json.load(open(f"gpurun_out/r3n_mpdata_shapes_{ns}.json"))
        for k,v in d["shapes"].items(): print(ns, k, v.get("ms_per_step"), v.get("frac_of_hbm"), v.get("launch"), v.get("bitwise_equal_to_oracle"), v.get("error"))
    except Exception as e: print(ns, "no result", e)
PY
tail -3 gpurun_out/r3n_mpdata_shapes.err
