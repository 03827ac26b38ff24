#!/bin/bash
# round 2, persistent strip sweep: parity suite, unit-length sweep at ne120, DRAM bytes of the chosen form
# usage (under gpurun): bash profiles/gpu_r2b.sh <tag>
tag=$1
mkdir -p gpurun_out
( time timeout 600 python -m pytest tests -m gpu -x -q ) > gpurun_out/${tag}_pytest.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_pytest.log
tail -15 gpurun_out/${tag}_pytest.log
for T in 0 3 4 6 9; do
  if [ $T = 0 ]; then export BWK_PERSIST=0; else export BWK_PERSIST=1 BWK_PERSIST_T=$T; fi
  echo "== persist T=$T"; timeout 300 python profiles/prof_run.py 120 72 40 2>&1 | grep -E "rep [234] sweep2|rror"
done
export BWK_PERSIST=1
for T in 4; do
BWK_PERSIST_T=$T timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"sweep2" -c 4 --csv \
  --log-file gpurun_out/${tag}_dram_ne120_T$T.csv python profiles/prof_run.py 120 72 40 > gpurun_out/${tag}_ncu_dram_ne120.log 2>&1
grep -E "sweep2" gpurun_out/${tag}_dram_ne120_T$T.csv | cut -d, -f5,9,13- | tail -12
done
