#!/bin/bash
# 8-GPU box: strong scaling at the shipped defaults (N = 8, 4, 2), group-size A/B at N = 4 and 8, multi-rank tests
tag=$1
mkdir -p gpurun_out
run() { n=$1; port=$2; shift 2; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n "$@"; }
( time timeout 900 python -m pytest tests/test_dist.py -m gpu -x -q ) > gpurun_out/${tag}_pytest_dist.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_pytest_dist.log
tail -5 gpurun_out/${tag}_pytest_dist.log | head -2
for n in 8 4; do
for cfg in BWK_WSTRIP_GROUP_MIN=4096 BWK_WSTRIP_GROUP_MIN=0 BWK_WSTRIP_GROUP_MIN=1000000; do
  env $cfg bash -c "$(declare -f run); run $n 29531 --steps 20 --warmup 6 --no-e2e --no-hv" 2>/dev/null | grep '^{' > gpurun_out/${tag}_ab.json
  python -c "import json; d=json.load(open('gpurun_out/${tag}_ab.json')); print('$cfg n=$n ms_per_step', round(d['ms_per_step'],4), d['checksum']['match'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"
done
done
for n in 8 4 2; do
  run $n 29517 --steps 20 --warmup 6 2> gpurun_out/${tag}_bench_n${n}.err | grep '^{' > gpurun_out/${tag}_bench_n${n}.json
  python -c "import json; d=json.load(open('gpurun_out/${tag}_bench_n${n}.json')); print('n=$n ms_per_step', round(d['ms_per_step'],4), d['checksum']['bits'], d['checksum']['match'], 'hv', d['advance_hypervis_scalar']['ms_per_step'], 'e2e', d['e2e']['value'], d['clocks']['sm_mhz'])"
done
