#!/bin/bash
# N-GPU check of the three-stream schedule: multi-rank tests, bench at N with the old and the new schedule
# usage (under gpurun --gpus N): bash profiles/gpu_r2n.sh <tag> <N>
tag=$1; n=$2
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_dist.py -m gpu -x -q ) > gpurun_out/${tag}_pytest_dist.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_pytest_dist.log
tail -6 gpurun_out/${tag}_pytest_dist.log
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $n "${@:2}"; }
for cfg in BWK_MULTI_SCHED=0 BWK_MULTI_SCHED=1 BWK_MULTI_SCHED=0 BWK_MULTI_SCHED=1; do
  env $cfg bash -c "$(declare -f run); n=$n; run 29531 --steps 20 --warmup 5 --no-e2e --no-hv" 2>/dev/null | grep '^{' > gpurun_out/${tag}_bench_${cfg}_n${n}.json
  python -c "import json; d=json.load(open('gpurun_out/${tag}_bench_${cfg}_n${n}.json')); print('$cfg n=$n ms_per_step', round(d['ms_per_step'],4), d['checksum']['bits'], d['checksum']['match'], 'lap', round(d['laplace_only']['ms_per_step'],4), d['clocks'])"
done
