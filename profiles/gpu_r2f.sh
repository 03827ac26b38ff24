#!/bin/bash
# ncu --set full (with source) of the warp-strip sweep and the element-local sweep 1 at ne30 L72 q40
tag=$1
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"wstrip|sweep_kernel" -s 4 -c 3 -o gpurun_out/${tag}_full_ne30 \
  python profiles/prof_run.py 30 72 40 > gpurun_out/${tag}_ncu_full.log 2>&1
tail -3 gpurun_out/${tag}_ncu_full.log; ls -la gpurun_out/${tag}_full_ne30.ncu-rep
