#!/bin/bash
# round 2, GPU call U (8 GPUs): C4 weak scaling, native mode (hypre's distributed setup, -P 2 2 2), 128^3 per GPU
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29519 --nproc-per-node 8 bench.py --gpus 8 --native --weak --workload lap27_128_gs11 --steps 5 --warmup 3 --no-aux > gpurun_out/r2u_bench_c4_128_native_8gpu.json 2> gpurun_out/r2u_bench_c4_128_native_8gpu.err
grep "^{" gpurun_out/r2u_bench_c4_128_native_8gpu.json | cut -c1-400; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/r2u_bench_c4_128_native_8gpu.err | head -5 | cut -c1-200
