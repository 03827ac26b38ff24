#!/bin/bash
# one GPU round-trip: parity tests, bench line, ncu launch list of the bench command, ncu --set full of the sweep kernels at ne30
# usage (under gpurun): bash profiles/gpu_round.sh <tag>
tag=$1
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/${tag}_pytest.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_pytest.log
tail -5 gpurun_out/${tag}_pytest.log
( time python bench.py ) > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc $?"
cat gpurun_out/${tag}_bench.json; tail -3 gpurun_out/${tag}_bench.err
python profiles/prof_run.py 120 72 40 > gpurun_out/${tag}_prof_ne120.log 2>&1; cat gpurun_out/${tag}_prof_ne120.log
python profiles/prof_run.py 30 72 40 > gpurun_out/${tag}_prof_ne30.log 2>&1; cat gpurun_out/${tag}_prof_ne30.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file gpurun_out/${tag}_launches_bench_ne120.csv python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/${tag}_ncu_launches.log 2>&1
tail -2 gpurun_out/${tag}_ncu_launches.log
ncu --set full --clock-control none --import-source on -k regex:sweep -s 3 -c 3 -o gpurun_out/${tag}_full_ne30 \
  python profiles/prof_run.py 30 72 40 > gpurun_out/${tag}_ncu_full.log 2>&1
tail -2 gpurun_out/${tag}_ncu_full.log
ls -la gpurun_out
# the caller (advance_hypervis_scalar): timing at ne120, ncu --set full of its three launches at ne30
python profiles/prof_hv.py > gpurun_out/${tag}_prof_hv_ne120.log 2>&1; tail -3 gpurun_out/${tag}_prof_hv_ne120.log
ncu --set full --clock-control none --import-source on -k regex:sweep -s 6 -c 3 -o gpurun_out/${tag}_full_ne30_hv \
  python profiles/prof_hv.py 30 72 40 > gpurun_out/${tag}_ncu_full_hv.log 2>&1
tail -2 gpurun_out/${tag}_ncu_full_hv.log
