#!/bin/bash
tag=$1
mkdir -p gpurun_out
for cfg in BWK_SWEEP2_TPU=5 BWK_SWEEP2_TPU=7 BWK_SWEEP2_TPU=9 BWK_SWEEP2_TPU=12 BWK_SWEEP2_TPU=15; do
  env $(echo $cfg | tr ',' ' ') timeout 300 python profiles/prof_run.py 120 72 40 2>&1 | grep -E "rep [234] sweep2|rror" | sed "s/^/[$cfg] /"
  env $(echo $cfg | tr ',' ' ') timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum --clock-control none -k regex:"sweep2_wstrip" -c 1 --csv \
    --log-file gpurun_out/${tag}_dram_$cfg.csv python profiles/prof_run.py 120 72 40 > gpurun_out/${tag}_ncu.log 2>&1
  grep -E "sweep2" gpurun_out/${tag}_dram_$cfg.csv | awk -F'","' '{print $9, $(NF-2), $NF}' | sed "s/^/[$cfg] /"
done
