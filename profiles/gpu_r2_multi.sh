#!/bin/bash
# round 2, multi-GPU round-trip (charged N x): the multi-rank GPU tests (log kept under profiles/), the per-phase trace of one
# step (BWK_TRACE serialises the phases), bench.py at N ranks incl. the pcie probe, strip-sweep form A/B at N
# usage (under gpurun --gpus N): bash profiles/gpu_r2_multi.sh <tag> <N> [more N ...]
tag=$1; shift
mkdir -p gpurun_out
nvidia-smi -L | head -8; nvidia-smi topo -m 2>/dev/null | head -12; nproc; free -g | head -2
( time timeout 900 python -m pytest tests/test_dist.py -m gpu -x -q -v ) > gpurun_out/${tag}_pytest_dist.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_pytest_dist.log
tail -12 gpurun_out/${tag}_pytest_dist.log
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 bench.py --gpus $1 "${@:3}"; }
for n in "$@"; do
  BWK_TRACE=1 run $n 29519 --steps 2 --warmup 1 --no-e2e --no-hv > gpurun_out/${tag}_trace_n${n}.json 2> gpurun_out/${tag}_trace_n${n}.txt
  grep "rank 0\]" gpurun_out/${tag}_trace_n${n}.txt | tail -6
  run $n 29517 --steps 10 --warmup 3 > gpurun_out/${tag}_bench_n${n}.json 2> gpurun_out/${tag}_bench_n${n}.err
  echo "bench n=$n rc $?"; cat gpurun_out/${tag}_bench_n${n}.json; tail -3 gpurun_out/${tag}_bench_n${n}.err
  BWK_STRIP_FORM=0 run $n 29521 --steps 10 --warmup 3 --no-e2e --no-hv > gpurun_out/${tag}_bench_form0_n${n}.json 2> /dev/null
  python -c "import json,sys; d=json.load(open('gpurun_out/${tag}_bench_form0_n${n}.json')); print('form0 n=$n ms_per_step', d['ms_per_step'], d['checksum']['bits'])"
done
