#!/bin/bash
# round 2, GPU call T (2 GPUs): final tree -- C3 strong and C4 weak (native) at 2 ranks
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29518"
timeout 600 $TR --nproc-per-node 2 bench.py --gpus 2 --steps 5 --warmup 3 --no-aux > gpurun_out/r2t_bench_c3_2gpu.json 2> gpurun_out/r2t_bench_c3_2gpu.err; grep "^{" gpurun_out/r2t_bench_c3_2gpu.json | cut -c1-260
timeout 600 $TR --nproc-per-node 2 bench.py --gpus 2 --native --weak --workload lap27_128_gs11 --steps 5 --warmup 3 --no-aux > gpurun_out/r2t_bench_c4_128_native_2gpu.json 2> gpurun_out/r2t_bench_c4_128_native_2gpu.err; grep "^{" gpurun_out/r2t_bench_c4_128_native_2gpu.json | cut -c1-400
