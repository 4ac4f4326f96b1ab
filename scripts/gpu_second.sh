#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python scripts/prof_ops.py > gpurun_out/ops_lap7_256.json 2> gpurun_out/ops_lap7_256.err
python scripts/prof_ops.py --gen laplacian_27pt --n 200 > gpurun_out/ops_lap27_200.json 2> gpurun_out/ops_lap27_200.err
for lanes in 1 2 4 8; do python scripts/prof_ops.py --layout 0 --lanes $lanes > gpurun_out/ops_lap7_256_csr$lanes.json 2>/dev/null; done
for lanes in 8 16 32; do python scripts/prof_ops.py --gen laplacian_27pt --n 200 --layout 0 --lanes $lanes > gpurun_out/ops_lap27_200_csr$lanes.json 2>/dev/null; done
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
python scripts/prof_ops.py --reps 2 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rows_kernel -s 3 -c 4 -o gpurun_out/prof_rows python scripts/prof_ops.py --reps 2 > gpurun_out/ncu_full.log 2>&1
tail -n 3 gpurun_out/pytest_gpu.log; cat gpurun_out/ops_*.json; tail -n 2 gpurun_out/ncu_launches.log gpurun_out/ncu_full.log
