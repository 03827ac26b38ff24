#!/bin/bash
# warp-strip sweep v2 (strip ends by cp.async): suite, timing at ne120 vs the round-1 form, DRAM bytes
tag=$1
mkdir -p gpurun_out
( time timeout 600 python -m pytest tests -m gpu -x -q ) > gpurun_out/${tag}_pytest.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_pytest.log
tail -6 gpurun_out/${tag}_pytest.log
for cfg in BWK_STRIP_FORM=0 BWK_STRIP_FORM=2 BWK_STRIP_FORM=2,BWK_SWEEP2_TPU=6 BWK_STRIP_FORM=2,BWK_SWEEP2_TPU=9 BWK_STRIP_FORM=2,BWK_SWEEP2_TPU=15 BWK_STRIP_FORM=0 BWK_STRIP_FORM=2; do
  env $(echo $cfg | tr ',' ' ') timeout 300 python profiles/prof_run.py 120 72 40 2>&1 | grep -E "rep [234] sweep2|rror" | sed "s/^/[$cfg] /"
done
for cfg in BWK_STRIP_FORM=2,BWK_SWEEP2_TPU=15 BWK_STRIP_FORM=2,BWK_SWEEP2_TPU=9; do
  env $(echo $cfg | tr ',' ' ') timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -k regex:"sweep2" -c 2 --csv \
    --log-file gpurun_out/${tag}_dram.csv python profiles/prof_run.py 120 72 40 > gpurun_out/${tag}_ncu.log 2>&1
  grep -E "sweep2.*(read|duration|inst_exec)" gpurun_out/${tag}_dram.csv | awk -F'","' '{print $5, $(NF-2), $NF}' | sed "s/^/[$cfg] /"
done
