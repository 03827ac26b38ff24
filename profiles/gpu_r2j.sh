#!/bin/bash
# warp-strip sweep: 4 CTAs/SM shape (2-stage ring, 128 registers) against the 3 CTAs/SM shape
tag=$1
mkdir -p gpurun_out
( time BWK_WSTRIP_SHAPE=4 timeout 600 python -m pytest tests -m gpu -x -q ) > gpurun_out/${tag}_pytest.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_pytest.log
tail -6 gpurun_out/${tag}_pytest.log
for cfg in BWK_WSTRIP_SHAPE=3 BWK_WSTRIP_SHAPE=4 BWK_WSTRIP_SHAPE=4,BWK_SWEEP2_TPU=9 BWK_WSTRIP_SHAPE=4,BWK_SWEEP2_TPU=15 BWK_WSTRIP_SHAPE=4,BWK_SWEEP2_TPU=23 BWK_WSTRIP_SHAPE=3 BWK_WSTRIP_SHAPE=4; do
  env $(echo $cfg | tr ',' ' ') timeout 300 python profiles/prof_run.py 120 72 40 2>&1 | grep -E "rep [234] sweep2|rror" | sed "s/^/[$cfg] /"
done
for cfg in BWK_WSTRIP_SHAPE=4; do
  env $(echo $cfg | tr ',' ' ') timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:"sweep2" -c 2 --csv \
    --log-file gpurun_out/${tag}_dram.csv python profiles/prof_run.py 120 72 40 > gpurun_out/${tag}_ncu.log 2>&1
  grep -E "sweep2" gpurun_out/${tag}_dram.csv | awk -F'","' '{print $5, $(NF-2), $NF}' | sed "s/^/[$cfg] /"
done
