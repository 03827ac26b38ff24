#!/bin/bash
# 8-GPU box: BASELINE config 5 (ne512 L128 q10) parity test + bench line, then the ne120 headline at N=8
tag=$1
mkdir -p gpurun_out
( time python -m pytest tests/test_dist.py -m gpu -x -q -k ne512 ) > gpurun_out/${tag}_pytest_ne512.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_pytest_ne512.log
tail -6 gpurun_out/${tag}_pytest_ne512.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29547"
timeout 600 $TR bench.py --gpus 8 --steps 5 --warmup 3 --ne 512 --nlev 128 --qsize 10 --no-e2e --no-hv > gpurun_out/${tag}_bench_ne512_n8.json 2> gpurun_out/${tag}_bench_ne512_n8.err
echo "bench ne512 rc $?"; cat gpurun_out/${tag}_bench_ne512_n8.json | cut -c1-900
timeout 600 $TR bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/${tag}_bench_ne120_n8.json 2> gpurun_out/${tag}_bench_ne120_n8.err
echo "bench ne120 rc $?"; cat gpurun_out/${tag}_bench_ne120_n8.json | cut -c1-1200
