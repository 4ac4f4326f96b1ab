#!/bin/bash
# round 2, GPU call Q (1 GPU): ncu evidence -- launch list of the bench command, --set full of the level-0 SpMV (pattern storage)
mkdir -p gpurun_out
timeout 600 python bench.py --steps 1 --warmup 1 --no-aux --no-cpu-baseline > gpurun_out/r2q_plain_bench.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2q_launches.csv python bench.py --steps 1 --warmup 1 --no-aux --no-cpu-baseline > gpurun_out/r2q_ncu_bench.log 2>&1
tail -n 2 gpurun_out/r2q_ncu_bench.log | cut -c1-200
timeout 300 python scripts/prof_ops.py --n 256 --reps 2 > gpurun_out/r2q_plain_ops.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:rows_kernel -s 2 -c 4 -o gpurun_out/r2q_prof python scripts/prof_ops.py --n 256 --reps 2 > gpurun_out/r2q_ncu_ops.log 2>&1
tail -n 3 gpurun_out/r2q_ncu_ops.log | cut -c1-200; wc -l gpurun_out/r2q_launches.csv
