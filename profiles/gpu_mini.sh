#!/bin/bash
# Mini-app round (nested_loops + mmf-mpdata-tracer): parity tests, timing with the
# launch-shape A/B, launch lists, one ncu --set full capture of each hot kernel.
# usage (under gpurun, one GPU): bash profiles/gpu_mini.sh <tag>
tag=$1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_nested.py tests/test_zzz_mpdata.py -m gpu -x -q > gpurun_out/${tag}_mini_pytest.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_mini_pytest.log
grep -v "^$" gpurun_out/${tag}_mini_pytest.log | tail -15
# timing (never under a profiler)
timeout 300 python profiles/prof_nested.py > gpurun_out/${tag}_nested_prof.json 2> gpurun_out/${tag}_nested_prof.err; tail -c 1500 gpurun_out/${tag}_nested_prof.json
timeout 300 python profiles/prof_mpdata.py > gpurun_out/${tag}_mpdata_prof.json 2> gpurun_out/${tag}_mpdata_prof.err; tail -c 2500 gpurun_out/${tag}_mpdata_prof.json
timeout 300 python profiles/prof_mpdata.py --flush --slices 48,1184 > gpurun_out/${tag}_mpdata_prof_cold.json 2>> gpurun_out/${tag}_mpdata_prof.err
# launch lists, then the full captures (each after its command has exited 0 without ncu, above)
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"advect_kernel" -c 12 --csv \
  --log-file gpurun_out/${tag}_mpdata_launches.csv python profiles/prof_mpdata.py --slices 48,1184 --iters 5 > gpurun_out/${tag}_ncu_mpdata.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"advect_kernel" -s 10 -c 2 -o gpurun_out/${tag}_mpdata_full \
  python profiles/prof_mpdata.py --slices 48 --iters 5 >> gpurun_out/${tag}_ncu_mpdata.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"flux_kernel|mask_tracer" -c 8 --csv \
  --log-file gpurun_out/${tag}_nested_launches.csv python profiles/prof_nested.py --quick > gpurun_out/${tag}_ncu_nested.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"flux_kernel" -c 2 -o gpurun_out/${tag}_nested_full \
  python profiles/prof_nested.py --quick >> gpurun_out/${tag}_ncu_nested.log 2>&1
ls -la gpurun_out | tail -12
