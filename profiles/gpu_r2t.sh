#!/bin/bash
# MPDATA at a production slice count: ncu --set full with source of advect_kernel (9,472 slices, default launch shape)
tag=$1
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"advect_kernel" -s 12 -c 1 -o gpurun_out/${tag}_mpdata_full_9472 \
  python profiles/prof_mpdata.py --slices 9472 --iters 5 > gpurun_out/${tag}_ncu_mpdata.log 2>&1
tail -3 gpurun_out/${tag}_ncu_mpdata.log; ls -la gpurun_out/${tag}_mpdata_full_9472.ncu-rep
