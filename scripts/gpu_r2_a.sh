#!/bin/bash
# round 2, GPU call A: parity of the new storage formats, per-level operator table, first bench line
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2a_box.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2a_pytest.log 2>&1; tail -n 12 gpurun_out/r2a_pytest.log
timeout 900 python scripts/fmt_probe.py --n 256 --levels 0,1,2 > gpurun_out/r2a_fmt_probe.jsonl 2> gpurun_out/r2a_fmt_probe.err; tail -n 3 gpurun_out/r2a_fmt_probe.err
timeout 600 python bench.py --steps 5 --warmup 3 --no-aux > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; tail -n 3 gpurun_out/r2a_bench.err; cat gpurun_out/r2a_bench.json | head -c 1500
