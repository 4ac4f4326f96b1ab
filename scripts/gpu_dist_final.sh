#!/bin/bash
# N-GPU final pass: parity worker, bench (both workloads), reference arm under torchrun
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544"
timeout 900 $TR tests/dist_gpu_worker.py > gpurun_out/dist${N}.log 2>&1; echo "rc=$?" >> gpurun_out/dist${N}.log
timeout 1500 $TR bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench${N}_256.log 2>&1; echo "rc=$?" >> gpurun_out/bench${N}_256.log
timeout 600 $TR bench.py --gpus $N --workload lap7_100_l1jac --steps 3 --warmup 3 > gpurun_out/bench${N}_100.log 2>&1; echo "rc=$?" >> gpurun_out/bench${N}_100.log
if [ "$N" == "2" ]; then
timeout 600 $TR bench.py --impl reference --gpus $N --steps 1 --warmup 1 > gpurun_out/bench${N}_reference.log 2>&1; echo "rc=$?" >> gpurun_out/bench${N}_reference.log
fi
tail -n 12 gpurun_out/dist${N}.log; for f in gpurun_out/bench${N}_*.log; do echo $f; tail -n 2 $f | cut -c 1-400; done
