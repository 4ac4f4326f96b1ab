#!/bin/bash
# round 2, GPU call D: pattern kernel v3 (dictionary in the constant bank, exact entry counts, one-wave grid)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/r2d_pytest.log 2>&1; tail -n 4 gpurun_out/r2d_pytest.log
timeout 900 python scripts/fmt_probe.py --n 256 --levels 0,1 --configs pat8 > gpurun_out/r2d_fmt_probe.jsonl 2> gpurun_out/r2d_fmt_probe.err; tail -n 3 gpurun_out/r2d_fmt_probe.err
timeout 900 python bench.py --steps 5 --warmup 3 --no-aux > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; tail -n 3 gpurun_out/r2d_bench.err; head -c 400 gpurun_out/r2d_bench.json
