#!/bin/bash
# persistent strip sweep, knob A/B at ne120 with DRAM bytes (one ncu pass of light metrics per setting)
tag=$1; shift
mkdir -p gpurun_out
for cfg in "$@"; do
  env $(echo $cfg | tr ',' ' ') timeout 300 python profiles/prof_run.py 120 72 40 2>&1 | grep -E "rep [34] sweep2|rror" | sed "s/^/[$cfg] /"
  env $(echo $cfg | tr ',' ' ') timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"sweep2" -c 2 --csv \
    --log-file gpurun_out/${tag}_dram.csv python profiles/prof_run.py 120 72 40 > gpurun_out/${tag}_ncu.log 2>&1
  grep -E "sweep2.*(read|duration)" gpurun_out/${tag}_dram.csv | awk -F'","' '{print $5, $(NF-2), $NF}' | sed "s/^/[$cfg] /"
done
