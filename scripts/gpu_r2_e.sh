#!/bin/bash
# round 2, GPU call E: fused PCG iteration (direction inside s = A p, <r,z> inside the last smoothing launch), zpad default
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2e_pytest.log 2>&1; tail -n 15 gpurun_out/r2e_pytest.log | cut -c1-300
timeout 900 python scripts/fmt_probe.py --n 256 --levels 0 --configs pat8 > gpurun_out/r2e_fmt_probe.jsonl 2> gpurun_out/r2e_fmt_probe.err; tail -n 3 gpurun_out/r2e_fmt_probe.err
timeout 900 python bench.py --steps 5 --warmup 3 --no-aux > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err; tail -n 3 gpurun_out/r2e_bench.err; head -c 400 gpurun_out/r2e_bench.json; echo
B200_PCG_FUSE=0 timeout 900 python bench.py --steps 5 --warmup 3 --no-aux --no-cpu-baseline > gpurun_out/r2e_bench_nofuse.json 2> gpurun_out/r2e_bench_nofuse.err; head -c 400 gpurun_out/r2e_bench_nofuse.json; echo
B200_PROFILE=1 timeout 600 python bench.py --steps 2 --warmup 1 --no-aux --no-cpu-baseline > gpurun_out/r2e_bench_prof.json 2> gpurun_out/r2e_bench_prof.err; tail -n 30 gpurun_out/r2e_bench_prof.err
