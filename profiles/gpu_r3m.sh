#!/bin/bash
# r3m: the allocation-free entry point (bwk_biharmonic_wk_scalar_noalloc): its test alone in a fresh process, then the
# whole GPU suite with that file collected last
mkdir -p gpurun_out
timeout 60 python -m pytest tests/test_zzzz_no_alloc.py -x -q -s -m gpu > gpurun_out/r3m_noalloc_alone.txt 2>&1; echo "alone rc $?"
tail -6 gpurun_out/r3m_noalloc_alone.txt
timeout 200 python -m pytest tests -x -q -m gpu > gpurun_out/r3m_pytest_gpu.txt 2>&1; echo "suite rc $?"
tail -4 gpurun_out/r3m_pytest_gpu.txt
