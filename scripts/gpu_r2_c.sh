#!/bin/bash
# round 2, GPU call C: ncu --set full of the pattern-storage SpMV (level-0 operators of the 7-pt 256^3 matrix)
mkdir -p gpurun_out
timeout 300 python scripts/prof_ops.py --n 256 --reps 2 > gpurun_out/r2c_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:rows_kernel -s 2 -c 4 -o gpurun_out/r2c_prof python scripts/prof_ops.py --n 256 --reps 2 > gpurun_out/r2c_ncu.log 2>&1
tail -n 5 gpurun_out/r2c_plain.log | cut -c1-600; tail -n 5 gpurun_out/r2c_ncu.log | cut -c1-300
