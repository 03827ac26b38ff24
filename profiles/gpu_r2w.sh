#!/bin/bash
tag=$1
mkdir -p gpurun_out
( time timeout 600 python -m pytest tests -m gpu -x -q ) > gpurun_out/${tag}_pytest.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_pytest.log
tail -6 gpurun_out/${tag}_pytest.log
for cfg in BWK_WSTRIP_MEMBERS=8 BWK_WSTRIP_MEMBERS=4 BWK_WSTRIP_MEMBERS=8 BWK_STRIP_FORM=0; do
  env $(echo $cfg | tr ',' ' ') timeout 300 python profiles/prof_run.py 120 72 40 2>&1 | grep -E "rep [234] sweep2|rror" | sed "s/^/[$cfg] /"
done
for cfg in BWK_WSTRIP_MEMBERS=8; do
  env $(echo $cfg | tr ',' ' ') timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,smsp__inst_executed.sum --clock-control none -k regex:"sweep2" -c 2 --csv \
    --log-file gpurun_out/${tag}_dram_$cfg.csv python profiles/prof_run.py 120 72 40 > gpurun_out/${tag}_ncu.log 2>&1
  grep -E "sweep2" gpurun_out/${tag}_dram_$cfg.csv | awk -F'","' '{print $5, $9, $(NF-2), $NF}' | sed "s/^/[$cfg] /"
done
