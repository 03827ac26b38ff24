#!/bin/bash
# one GPU iteration: parity tests, kernel timings at ne30/ne120, then ONE ncu --set full capture of the sweep kernels at ne30
# usage (under gpurun): bash profiles/gpu_iter.sh <tag> [noncu]
tag=$1
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_pytest.log
tail -3 gpurun_out/${tag}_pytest.log
python profiles/prof_run.py 120 72 40 > gpurun_out/${tag}_prof_ne120.log 2>&1; cat gpurun_out/${tag}_prof_ne120.log
if [ "$2" != "noncu" ]; then
python profiles/prof_run.py 30 72 40 > gpurun_out/${tag}_prof_ne30.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:sweep -s 3 -c 3 -o gpurun_out/${tag}_full_ne30 python profiles/prof_run.py 30 72 40 > gpurun_out/${tag}_ncu_full.log 2>&1
cat gpurun_out/${tag}_prof_ne30.log; tail -2 gpurun_out/${tag}_ncu_full.log
fi
