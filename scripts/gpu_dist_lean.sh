#!/bin/bash
# N-GPU lean pass: parity worker + the 256^3 bench
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544"
timeout 300 $TR tests/dist_gpu_worker.py > gpurun_out/dist${N}.log 2>&1; echo "rc=$?" >> gpurun_out/dist${N}.log
timeout 400 $TR bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench${N}_256.log 2>&1; echo "rc=$?" >> gpurun_out/bench${N}_256.log
tail -n 9 gpurun_out/dist${N}.log; tail -n 2 gpurun_out/bench${N}_256.log | cut -c 1-1200
