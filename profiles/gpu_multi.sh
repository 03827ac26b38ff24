#!/bin/bash
# multi-GPU round-trip: the multi-rank GPU tests, then bench.py at N ranks (ours + reference arm)
# usage (under gpurun --gpus N): bash profiles/gpu_multi.sh <tag> <N> [more N ...]
tag=$1; shift
mkdir -p gpurun_out
nvidia-smi -L
( time python -m pytest tests -m gpu -x -q -k "multi or rank" ) > gpurun_out/${tag}_pytest_multi.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_pytest_multi.log
tail -5 gpurun_out/${tag}_pytest_multi.log
for n in "$@"; do
  if [ "$n" = "1" ]; then
    python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu > gpurun_out/${tag}_bench_n${n}.json 2> gpurun_out/${tag}_bench_n${n}.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 \
      bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/${tag}_bench_n${n}.json 2> gpurun_out/${tag}_bench_n${n}.err
  fi
  echo "bench n=$n rc $?"; cat gpurun_out/${tag}_bench_n${n}.json; tail -3 gpurun_out/${tag}_bench_n${n}.err
done
