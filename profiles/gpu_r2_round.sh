#!/bin/bash
# round 2, one-GPU evidence round at the shipped defaults: suite, bench (ours + CPU arm), launch list of the bench command,
# ne120 DRAM bytes per sweep kernel (one launch each, light metrics), ncu --set full of the sweep kernels at ne30
# usage (under gpurun): bash profiles/gpu_r2_round.sh <tag>
tag=$1
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/${tag}_pytest.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_pytest.log
tail -5 gpurun_out/${tag}_pytest.log
( time python bench.py --pcie ) > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc $?"
cat gpurun_out/${tag}_bench.json | cut -c1-3000; tail -3 gpurun_out/${tag}_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file gpurun_out/${tag}_launches_bench_ne120.csv python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --no-mini > gpurun_out/${tag}_ncu_launches.log 2>&1
tail -2 gpurun_out/${tag}_ncu_launches.log
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -k regex:"sweep" -s 4 -c 4 --csv \
  --log-file gpurun_out/${tag}_dram_ne120.csv python profiles/prof_run.py 120 72 40 > gpurun_out/${tag}_ncu_dram_ne120.log 2>&1
grep -E "sweep" gpurun_out/${tag}_dram_ne120.csv | awk -F'","' '{print $5, $9, $(NF-2), $NF}'
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"sweep" -s 4 -c 4 -o gpurun_out/${tag}_full_ne30 \
  python profiles/prof_run.py 30 72 40 > gpurun_out/${tag}_ncu_full.log 2>&1
tail -2 gpurun_out/${tag}_ncu_full.log
