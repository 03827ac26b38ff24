#!/bin/bash
# round 2, first GPU call (one GPU): the whole -m gpu suite incl. the new tests at the timed configs, the bench line (ours and the
# CPU arm at the headline size), the two mini-apps (timing, launch lists, ncu --set full), and a first-hand ne120 DRAM-bytes capture
# usage (under gpurun): bash profiles/gpu_r2a.sh <tag>
tag=$1
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv; free -g | head -2; nproc
( time python -m pytest tests -m gpu -x -q --durations=8 ) > gpurun_out/${tag}_pytest.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_pytest.log
tail -22 gpurun_out/${tag}_pytest.log
( time python bench.py --pcie ) > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc $?"
cat gpurun_out/${tag}_bench.json; tail -5 gpurun_out/${tag}_bench.err
( time python bench.py --impl reference --steps 3 --warmup 1 ) > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err; echo "ref rc $?"
cat gpurun_out/${tag}_bench_ref.json; tail -5 gpurun_out/${tag}_bench_ref.err
python profiles/prof_run.py 120 72 40 > gpurun_out/${tag}_prof_ne120.log 2>&1; tail -6 gpurun_out/${tag}_prof_ne120.log
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"sweep" -c 6 --csv \
  --log-file gpurun_out/${tag}_dram_ne120.csv python profiles/prof_run.py 120 72 40 > gpurun_out/${tag}_ncu_dram_ne120.log 2>&1
tail -8 gpurun_out/${tag}_dram_ne120.csv
bash profiles/gpu_mini.sh ${tag}
