#!/bin/bash
# round 2, GPU call L (1 GPU): the no-allocation dot entry point -- alone in a fresh process, then after the whole suite
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r2l_noalloc_alone.log 2>&1; tail -n 6 gpurun_out/r2l_noalloc_alone.log | cut -c1-300
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2l_pytest_all.log 2>&1; tail -n 8 gpurun_out/r2l_pytest_all.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2l_smoke.log 2>&1; tail -n 3 gpurun_out/r2l_smoke.log
