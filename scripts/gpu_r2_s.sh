#!/bin/bash
# round 2, GPU call S (1 GPU): final validation of the committed tree
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2s_pytest.log 2>&1; tail -n 6 gpurun_out/r2s_pytest.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2s_smoke.log 2>&1; tail -n 2 gpurun_out/r2s_smoke.log
timeout 900 python bench.py > gpurun_out/r2s_bench.json 2> gpurun_out/r2s_bench.err; tail -n 2 gpurun_out/r2s_bench.err | cut -c1-200; head -c 300 gpurun_out/r2s_bench.json; echo
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2s_bench_reference.json 2> gpurun_out/r2s_bench_reference.err; head -c 300 gpurun_out/r2s_bench_reference.json; echo
