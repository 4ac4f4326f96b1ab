#!/bin/bash
mkdir -p gpurun_out
free -g > gpurun_out/mem.txt
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544"
timeout 900 $TR tests/dist_gpu_worker.py > gpurun_out/dist2.log 2>&1; echo "rc=$?" >> gpurun_out/dist2.log
timeout 900 $TR bench.py --gpus 2 --workload lap7_100_l1jac --steps 3 --warmup 3 > gpurun_out/bench2_100.log 2>&1; echo "rc=$?" >> gpurun_out/bench2_100.log
timeout 1500 $TR bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench2_256.log 2>&1; echo "rc=$?" >> gpurun_out/bench2_256.log
tail -n 5 gpurun_out/pytest_gpu.log; tail -n 12 gpurun_out/dist2.log; tail -n 4 gpurun_out/bench2_100.log gpurun_out/bench2_256.log
