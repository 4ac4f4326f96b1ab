#!/bin/bash
# round 2, GPU call O (8 GPUs): C3 strong scaling (default and with level 4 distributed), C2 SpMV, C4 weak (native) at 8 ranks
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29517"
N=8
free -g | head -2 > gpurun_out/r2o_box.txt; nproc >> gpurun_out/r2o_box.txt; nvidia-smi -L >> gpurun_out/r2o_box.txt
timeout 900 $TR --nproc-per-node $N bench.py --gpus $N --steps 5 --warmup 3 --no-aux > gpurun_out/r2o_bench_c3_${N}gpu.json 2> gpurun_out/r2o_bench_c3_${N}gpu.err; grep "^{" gpurun_out/r2o_bench_c3_${N}gpu.json | cut -c1-250; tail -n 2 gpurun_out/r2o_bench_c3_${N}gpu.err | cut -c1-200
B200_REPL_NNZ=2e6 timeout 900 $TR --nproc-per-node $N bench.py --gpus $N --steps 5 --warmup 3 --no-aux > gpurun_out/r2o_bench_c3_${N}gpu_repl2e6.json 2> gpurun_out/r2o_bench_c3_${N}gpu_repl2e6.err; grep "^{" gpurun_out/r2o_bench_c3_${N}gpu_repl2e6.json | cut -c1-250
timeout 600 $TR --nproc-per-node $N scripts/spmv_bench.py > gpurun_out/r2o_spmv_c2_${N}gpu.json 2> gpurun_out/r2o_spmv_c2_${N}gpu.err; grep "^{" gpurun_out/r2o_spmv_c2_${N}gpu.json | cut -c1-400
timeout 900 $TR --nproc-per-node $N bench.py --gpus $N --native --weak --workload lap27_128_gs11 --steps 5 --warmup 3 --no-aux > gpurun_out/r2o_bench_c4_128_native_${N}gpu.json 2> gpurun_out/r2o_bench_c4_128_native_${N}gpu.err; grep "^{" gpurun_out/r2o_bench_c4_128_native_${N}gpu.json | cut -c1-400; tail -n 2 gpurun_out/r2o_bench_c4_128_native_${N}gpu.err | cut -c1-200
