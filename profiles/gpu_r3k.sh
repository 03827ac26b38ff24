#!/bin/bash
# r3k: MPDATA marching form (mpdata_march.cuh): GPU tests, timing against the barrier-phased kernel at 9,472 / 1,184 / 48 slices,
# one ncu --set full at 9,472 slices
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zzz_mpdata.py -m gpu -x -q > gpurun_out/r3k_pytest_mpdata.txt 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r3k_pytest_mpdata.txt
for ns in 9472 1184 48; do
  timeout 200 python profiles/prof_mpdata_shapes.py --slices $ns --shapes 0x0,m4,m6,m8,m12,m13 > gpurun_out/r3k_mpdata_shapes_$ns.json 2>> gpurun_out/r3k_mpdata_shapes.err; echo "shapes $ns rc $?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r3k_mpdata_shapes_$ns.json"))
for k,v in d["shapes"].items(): print($ns, k, v)
PY
done
timeout 200 ncu --set full --clock-control none --import-source on -k regex:advect_march -s 1 -c 1 -f -o gpurun_out/r3k_mpdata_march_9472 \
  python profiles/prof_mpdata_shapes.py --shapes m12 --iters 3 --no-check > gpurun_out/r3k_ncu.log 2>&1; echo "ncu rc $?"
