#!/bin/bash
# round 2, multi-GPU round-trip (charged N x): the multi-rank GPU tests (log kept under profiles/), the per-phase trace of one
# step (BWK_TRACE serialises the phases), bench.py at N ranks (ours incl. the pcie probe, and the CPU arm)
# usage (under gpurun --gpus N): bash profiles/gpu_r2_multi.sh <tag> <N> [more N ...]
tag=$1; shift
mkdir -p gpurun_out
nvidia-smi -L | head -8; nvidia-smi topo -m 2>/dev/null | head -14; numactl -H 2>/dev/null | head -6; nproc; free -g | head -2
( time timeout 900 python -m pytest tests/test_dist.py -m gpu -x -q -v ) > gpurun_out/${tag}_pytest_dist.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_pytest_dist.log
tail -12 gpurun_out/${tag}_pytest_dist.log
for n in "$@"; do
  BWK_TRACE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29519 \
    bench.py --gpus $n --steps 2 --warmup 1 --no-e2e --no-hv > gpurun_out/${tag}_trace_n${n}.json 2> gpurun_out/${tag}_trace_n${n}.txt
  grep "rank 0\]" gpurun_out/${tag}_trace_n${n}.txt | tail -7
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/${tag}_bench_n${n}.json 2> gpurun_out/${tag}_bench_n${n}.err
  echo "bench n=$n rc $?"; cat gpurun_out/${tag}_bench_n${n}.json; tail -3 gpurun_out/${tag}_bench_n${n}.err
done
