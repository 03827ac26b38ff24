#!/bin/bash
# r3i: sanity after moving the strip grouping into bwk_plan.cpp (host code only): GPU suite, smoke, one bench line
mkdir -p gpurun_out
timeout 240 python -m pytest tests -m gpu -x -q > gpurun_out/r3i_pytest_gpu.txt 2>&1; echo "pytest rc $?"
tail -3 gpurun_out/r3i_pytest_gpu.txt
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r3i_smoke.txt 2>&1; echo "smoke rc $?"; tail -2 gpurun_out/r3i_smoke.txt
timeout 300 python bench.py --steps 10 --warmup 3 2> gpurun_out/r3i_bench.err | grep '^{' > gpurun_out/r3i_bench.json; echo "bench rc $?"
python -c "import json;d=json.load(open('gpurun_out/r3i_bench.json'));print(d['ms_per_step'],d['value'],d.get('checksum'),d['roofline']['frac'],d['e2e']['value'],d['clocks'])"
