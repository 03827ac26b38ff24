#!/bin/bash
# warp-strip sweep with 4 / 8 / 16 members per warp: suite at 8 and 16, timing and DRAM bytes at ne120
tag=$1
mkdir -p gpurun_out
for m in 8 16; do
( time BWK_WSTRIP_MEMBERS=$m timeout 600 python -m pytest tests -m gpu -x -q ) > gpurun_out/${tag}_pytest_m$m.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_pytest_m$m.log
tail -4 gpurun_out/${tag}_pytest_m$m.log | head -2
done
for cfg in BWK_WSTRIP_MEMBERS=4 BWK_WSTRIP_MEMBERS=8 BWK_WSTRIP_MEMBERS=16 BWK_WSTRIP_MEMBERS=8,BWK_SWEEP2_TPU=9 BWK_WSTRIP_MEMBERS=16,BWK_SWEEP2_TPU=9 BWK_WSTRIP_MEMBERS=4 BWK_WSTRIP_MEMBERS=8 BWK_WSTRIP_MEMBERS=16; do
  env $(echo $cfg | tr ',' ' ') timeout 300 python profiles/prof_run.py 120 72 40 2>&1 | grep -E "rep [234] sweep2|rror" | sed "s/^/[$cfg] /"
done
for cfg in BWK_WSTRIP_MEMBERS=8 BWK_WSTRIP_MEMBERS=16; do
  env $(echo $cfg | tr ',' ' ') timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,smsp__inst_executed.sum --clock-control none -k regex:"sweep2" -c 2 --csv \
    --log-file gpurun_out/${tag}_dram_$cfg.csv python profiles/prof_run.py 120 72 40 > gpurun_out/${tag}_ncu.log 2>&1
  grep -E "sweep2" gpurun_out/${tag}_dram_$cfg.csv | awk -F'","' '{print $5, $9, $(NF-2), $NF}' | sed "s/^/[$cfg] /"
done
