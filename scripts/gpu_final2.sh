#!/bin/bash
# final single-GPU evidence (compressed layout): bench line, ncu full capture of the row kernel, launch list of one solve
mkdir -p gpurun_out
python bench.py > gpurun_out/bench1_256_default.log 2>&1; tail -n 1 gpurun_out/bench1_256_default.log | cut -c 1-600
python scripts/prof_ops.py --reps 2 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rows_kernel -s 3 -c 4 -f -o gpurun_out/prof_rows_c python scripts/prof_ops.py --reps 2 > gpurun_out/ncu_full.log 2>&1
python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1300 --csv --log-file gpurun_out/launches_c.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
tail -n 2 gpurun_out/ncu_full.log; tail -n 1 gpurun_out/ncu_launches.log | cut -c 1-200
