#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29545"
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -n 3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --workload lap7_100_l1jac --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench1_100.log 2>&1
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench1_256.log 2>&1
bash scripts/gpu_dist.sh 2
B200_PROFILE=1 timeout 900 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/prof1_256.log 2>&1
B200_PROFILE=1 timeout 900 $TR bench.py --gpus 2 --steps 2 --warmup 1 > gpurun_out/prof2_256.log 2>&1
for f in gpurun_out/bench1_*.log; do echo $f; tail -n 1 $f | cut -c 1-300; done
