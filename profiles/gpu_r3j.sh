#!/bin/bash
# r3j: MPDATA at 9,472 slices -- launch-shape A/B of the shipped kernel (3 and 4 CTAs per SM), then one ncu --set full of the default shape
mkdir -p gpurun_out
timeout 200 python profiles/prof_mpdata_shapes.py > gpurun_out/r3j_mpdata_shapes.json 2> gpurun_out/r3j_mpdata_shapes.err; echo "shapes rc $?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r3j_mpdata_shapes.json"))
for k,v in d["shapes"].items(): print(k, v)
PY
timeout 200 ncu --set full --clock-control none --import-source on -k regex:advect_kernel -s 1 -c 1 -f -o gpurun_out/r3j_mpdata_9472 \
  python profiles/prof_mpdata_shapes.py --shapes 0x0 --iters 3 --no-check > gpurun_out/r3j_ncu.log 2>&1; echo "ncu rc $?"
ls -la gpurun_out/ | tail -5
