#!/bin/bash
# r3o: where the marching kernel (6 warps per CTA) overtakes the barrier-phased one: both forms explicitly and the automatic
# choice (0x0) at 1,776 ... 9,472 slices, each result compared with the oracle bit for bit
mkdir -p gpurun_out
for ns in 1776 2368 3552 4736 7104 9472; do
  timeout 12 python profiles/prof_mpdata_shapes.py --slices $ns --shapes 2x512,m6,0x0 > gpurun_out/r3o_mpdata_shapes_$ns.json 2>> gpurun_out/r3o_mpdata_shapes.err; echo "shapes $ns rc $?"
done
python - <<'PY'
import json
for ns in (1776, 2368, 3552, 4736, 7104, 9472):
    try:
        d=json.load(open(f"gpurun_out/r3o_mpdata_shapes_{ns}.json"))
        for k,v in d["shapes"].items(): print(ns, k, v.get("ms_per_step"), v.get("launch"), v.get("bitwise_equal_to_oracle"), v.get("error"))
    except Exception as e: print(ns, "no result", e)
PY
tail -3 gpurun_out/r3o_mpdata_shapes.err
