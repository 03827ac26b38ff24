#!/bin/bash
# 8-GPU confirmation of the whole-strip split + three-stream schedule: trace, schedule A/B, the full bench line at N = 8 and 4
tag=$1
mkdir -p gpurun_out
run() { n=$1; port=$2; shift 2; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n "$@"; }
BWK_TRACE=1 run 8 29519 --steps 2 --warmup 1 --no-e2e --no-hv > /dev/null 2> gpurun_out/${tag}_trace_n8.txt
grep "rank 0\]" gpurun_out/${tag}_trace_n8.txt | tail -6
for cfg in BWK_MULTI_SCHED=0 BWK_MULTI_SCHED=1 BWK_MULTI_SCHED=0 BWK_MULTI_SCHED=1; do
  env $cfg bash -c "$(declare -f run); run 8 29531 --steps 20 --warmup 5 --no-e2e --no-hv" 2>/dev/null | grep '^{' > gpurun_out/${tag}_bench_${cfg}_n8.json
  python -c "import json; d=json.load(open('gpurun_out/${tag}_bench_${cfg}_n8.json')); print('$cfg n=8 ms_per_step', round(d['ms_per_step'],4), d['checksum']['bits'], d['checksum']['match'], 'lap', round(d['laplace_only']['ms_per_step'],4), d['clocks'])"
done
for n in 8 4; do
  run $n 29517 --steps 20 --warmup 5 2> gpurun_out/${tag}_bench_n${n}.err | grep '^{' > gpurun_out/${tag}_bench_n${n}.json
  python -c "import json; d=json.load(open('gpurun_out/${tag}_bench_n${n}.json')); print('n=$n ms_per_step', round(d['ms_per_step'],4), d['checksum']['bits'], d['checksum']['match'], 'hv', d['advance_hypervis_scalar']['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['GBps_per_rank_per_direction'], d['e2e']['pcie']['both_GBps_per_rank_per_direction'])"
done
