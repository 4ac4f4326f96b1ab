#!/bin/bash
# round 2, one GPU: ncu evidence.  Each ncu pass runs only after the same command exited 0 without ncu (&&).
#   launch list (gpu__time_duration.sum) of the bench command      -> profiles/r02_launches_lap7_256_cheby.csv (+ _summary.txt
#                                                                       by scripts/summarize_launches.py)
#   --set full of the level-0 SpMV in pattern storage              -> gpurun_out/r02_prof.ncu-rep, read here with
#       ncu -i ... --page details > profiles/r02_ncu_full_rows_kernel_pat8.txt;  scripts/ncu_traffic.py -> r02_ncu_traffic_pat8.json
mkdir -p gpurun_out; O=gpurun_out/r02
timeout 600 python bench.py --steps 1 --warmup 1 --no-aux --no-cpu-baseline > ${O}_plain_bench.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file ${O}_launches.csv python bench.py --steps 1 --warmup 1 --no-aux --no-cpu-baseline > ${O}_ncu_bench.log 2>&1
tail -n 2 ${O}_ncu_bench.log | cut -c1-200
timeout 300 python scripts/prof_ops.py --n 256 --reps 2 > ${O}_plain_ops.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:rows_kernel -s 2 -c 4 -o ${O}_prof python scripts/prof_ops.py --n 256 --reps 2 > ${O}_ncu_ops.log 2>&1
tail -n 3 ${O}_ncu_ops.log | cut -c1-200; wc -l ${O}_launches.csv
