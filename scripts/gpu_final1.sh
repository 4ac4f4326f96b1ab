#!/bin/bash
# final single-GPU evidence: tests, bench line, ncu full capture of the row kernel, ncu launch list of one solve
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; tail -n 3 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench1_256_default.log 2>&1; tail -n 1 gpurun_out/bench1_256_default.log | cut -c 1-400
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.log 2>&1; tail -n 1 gpurun_out/bench_reference.log | cut -c 1-300
python bench.py --workload lap7_100_l1jac --steps 5 --warmup 3 > gpurun_out/bench1_100.log 2>&1; tail -n 1 gpurun_out/bench1_100.log | cut -c 1-300
python bench.py --workload lap27_128_gs11 --steps 5 --warmup 3 > gpurun_out/bench1_gs11.log 2>&1; tail -n 1 gpurun_out/bench1_gs11.log | cut -c 1-300
python scripts/prof_ops.py --reps 2 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rows_kernel -s 3 -c 4 -o gpurun_out/prof_rows python scripts/prof_ops.py --reps 2 > gpurun_out/ncu_full.log 2>&1
python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1300 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
python scripts/prof_ops.py > gpurun_out/ops_lap7_256.json 2>/dev/null; python scripts/prof_ops.py --gen laplacian_27pt --n 200 > gpurun_out/ops_lap27_200.json 2>/dev/null
tail -n 2 gpurun_out/ncu_full.log
