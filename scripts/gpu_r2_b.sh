#!/bin/bash
# round 2, GPU call B: pattern kernel with fixed-stride headers + header prefetch; per-kernel breakdown of the solve
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/r2b_pytest.log 2>&1; tail -n 4 gpurun_out/r2b_pytest.log
timeout 900 python scripts/fmt_probe.py --n 256 --levels 0,1,2 --configs pat8,auto_zpad > gpurun_out/r2b_fmt_probe.jsonl 2> gpurun_out/r2b_fmt_probe.err; tail -n 3 gpurun_out/r2b_fmt_probe.err
timeout 600 python bench.py --steps 5 --warmup 3 --no-aux > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; tail -n 3 gpurun_out/r2b_bench.err; head -c 600 gpurun_out/r2b_bench.json
B200_PROFILE=1 timeout 600 python bench.py --steps 2 --warmup 1 --no-aux --no-cpu-baseline > gpurun_out/r2b_bench_prof.json 2> gpurun_out/r2b_bench_prof.err; tail -n 60 gpurun_out/r2b_bench_prof.err
