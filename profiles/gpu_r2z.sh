#!/bin/bash
tag=$1
mkdir -p gpurun_out
( time timeout 600 python -m pytest tests -m gpu -x -q ) > gpurun_out/${tag}_pytest.log 2>&1; echo "rc $?" >> gpurun_out/${tag}_pytest.log
tail -6 gpurun_out/${tag}_pytest.log | head -3
for cfg in BWK_X=0 BWK_X=0 BWK_WSTRIP_MEMBERS=4; do
  env $(echo $cfg | tr ',' ' ') timeout 300 python profiles/prof_run.py 120 72 40 2>&1 | grep -E "rep [1234] sweep2|rror" | sed "s/^/[$cfg] /"
done
for cfg in BWK_WSTRIP_GROUP_MIN=4096 BWK_WSTRIP_GROUP_MIN=0; do
  echo "== fake rank 0 of 8, $cfg"
  env BWK_FAKE_RANKS=0,8 BWK_TRACE=1 $cfg timeout 300 python profiles/prof_fake_rank.py 120 72 40 6 2>&1 | grep -A7 "rep 5" | grep -E "produce|consume|strips"
done
python bench.py --steps 10 --warmup 5 --no-e2e --no-cpu --no-mini --no-hv --no-dp3d | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('bench ms', d['ms_per_step'], d['checksum']['match'], d['clocks'], d['roofline']['kernels'])"
