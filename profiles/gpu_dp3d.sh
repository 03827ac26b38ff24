#!/bin/bash
# dp3d iteration: its parity tests, timing at ne120 L72, launch durations of its three kernels
# usage (under gpurun): bash profiles/gpu_dp3d.sh <tag> [all]
tag=$1
mkdir -p gpurun_out
sel="tests/test_dp3d.py"; [ "$2" == "all" ] && sel="tests"
timeout 600 python -m pytest $sel -m gpu -x -q -s > gpurun_out/${tag}_pytest.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_pytest.log
grep -v "^$" gpurun_out/${tag}_pytest.log | tail -25
timeout 300 python profiles/prof_dp3d.py 120 72 > gpurun_out/${tag}_prof_dp3d_ne120.log 2>&1; cat gpurun_out/${tag}_prof_dp3d_ne120.log | tail -8
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"dp3d|strip" -c 6 --csv --log-file gpurun_out/${tag}_dp3d_launches_ne30.csv python profiles/prof_dp3d.py 30 72 > gpurun_out/${tag}_ncu_dp3d.log 2>&1
tail -20 gpurun_out/${tag}_dp3d_launches_ne30.csv | cut -c1-260
