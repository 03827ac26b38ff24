#!/bin/bash
# r3l: MPDATA marching forms after the warp-uniform sum loops: one slice per warp (m<warps>) and two (d<warps>); GPU tests,
# timing at 9,472 / 1,184 slices, ncu --set full of the better of m6 / d6 at 9,472 slices
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zzz_mpdata.py -m gpu -x -q > gpurun_out/r3l_pytest_mpdata.txt 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r3l_pytest_mpdata.txt
for ns in 9472 1184; do
  timeout 200 python profiles/prof_mpdata_shapes.py --slices $ns --shapes 0x0,m4,m6,m12,d2,d3,d6 > gpurun_out/r3l_mpdata_shapes_$ns.json 2>> gpurun_out/r3l_mpdata_shapes.err; echo "shapes $ns rc $?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r3l_mpdata_shapes_$ns.json"))
for k,v in d["shapes"].items(): print($ns, k, v.get("ms_per_step"), v.get("launch"), v.get("bitwise_equal_to_oracle"), v.get("error"))
PY
done
for sh in m6 d6; do
timeout 200 ncu --set full --clock-control none --import-source on -k regex:advect_march -s 1 -c 1 -f -o gpurun_out/r3l_mpdata_${sh}_9472 \
  python profiles/prof_mpdata_shapes.py --shapes $sh --iters 3 --no-check > gpurun_out/r3l_ncu_$sh.log 2>&1; echo "ncu $sh rc $?"
done
