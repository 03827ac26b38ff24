#!/bin/bash
# where does the warp-strip sweep's time go at ne120?  timing with the neighbour traffic switched off (results wrong on purpose)
tag=$1
mkdir -p gpurun_out
nvidia-smi --query-gpu=clocks.sm,power.draw,power.limit --format=csv,noheader
for cfg in BWK_DEBUG_FLAGS=0 BWK_DEBUG_FLAGS=1 BWK_DEBUG_FLAGS=2 BWK_DEBUG_FLAGS=3 BWK_DEBUG_FLAGS=0; do
  env $(echo $cfg | tr ',' ' ') timeout 300 python profiles/prof_run.py 120 72 40 2>&1 | grep -E "rep [234] sweep2|rror" | sed "s/^/[$cfg] /"
done
for cfg in BWK_DEBUG_FLAGS=3; do
  env $(echo $cfg | tr ',' ' ') timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -k regex:"sweep2" -c 2 --csv \
    --log-file gpurun_out/${tag}_dram.csv python profiles/prof_run.py 120 72 40 > gpurun_out/${tag}_ncu.log 2>&1
  grep -E "sweep2.*(read|duration|inst_exec)" gpurun_out/${tag}_dram.csv | awk -F'","' '{print $5, $(NF-2), $NF}' | sed "s/^/[$cfg] /"
done
