#!/bin/bash
# one rank of 8 on one GPU: per-phase cost by strip-sweep form and unit length
tag=$1
mkdir -p gpurun_out
for cfg in BWK_STRIP_FORM=2 BWK_STRIP_FORM=0 BWK_STRIP_FORM=2,BWK_SWEEP2_TPU=9 BWK_STRIP_FORM=2,BWK_SWEEP2_TPU=15 BWK_STRIP_FORM=2,BWK_SWEEP2_TPU=23 BWK_STRIP_FORM=2,BWK_SWEEP2_TPU=30; do
  echo "== $cfg"
  env BWK_FAKE_RANKS=0,8 BWK_TRACE=1 $(echo $cfg | tr ',' ' ') timeout 300 python profiles/prof_fake_rank.py 120 72 40 4 2>&1 | grep -A7 "rep 3" | grep -E "produce|consume|strips"
done
echo "== full sphere for reference (same box)"; timeout 300 python profiles/prof_run.py 120 72 40 2>&1 | grep -E "rep 4"
BWK_FAKE_RANKS=0,8 BWK_TRACE=1 timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"sweep" -c 12 --csv \
    --log-file gpurun_out/${tag}_dram.csv python profiles/prof_fake_rank.py 120 72 40 2 > gpurun_out/${tag}_ncu.log 2>&1
grep -E "sweep.*(read|write|duration)" gpurun_out/${tag}_dram.csv | awk -F'","' '{print $5, $9, $(NF-2), $NF}' | tail -18
