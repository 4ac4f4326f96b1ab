#!/bin/bash
# last check of the committed tree: bench line (with the aux microbench), smoke(), GPU tests
mkdir -p gpurun_out
python bench.py > gpurun_out/bench1_256_final.log 2>&1; tail -n 1 gpurun_out/bench1_256_final.log | cut -c 1-700
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; tail -n 3 gpurun_out/smoke_final.log
timeout 70 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_final.log 2>&1; tail -n 3 gpurun_out/pytest_gpu_final.log
