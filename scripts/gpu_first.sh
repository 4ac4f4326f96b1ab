#!/bin/bash
# first GPU pass: tests, smoke, small + full bench (logs under gpurun_out/)
mkdir -p gpurun_out
nvidia-smi > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/host.txt; lscpu | head -20 >> gpurun_out/host.txt; which mpirun mpicc >> gpurun_out/host.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --workload lap7_100_l1jac --steps 3 --warmup 3 > gpurun_out/bench_100.log 2>&1; echo "rc=$?" >> gpurun_out/bench_100.log
timeout 1500 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_256.log 2>&1; echo "rc=$?" >> gpurun_out/bench_256.log
tail -3 gpurun_out/pytest_gpu.log gpurun_out/smoke.log gpurun_out/bench_100.log gpurun_out/bench_256.log
