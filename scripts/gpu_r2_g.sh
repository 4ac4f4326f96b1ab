#!/bin/bash
# round 2, GPU call G (2 GPUs): multi-rank parity tests, C2 SpMV microbench at 1 and 2 GPUs, C3 at 2 GPUs with per-kernel profile
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29511"
timeout 900 python -m pytest tests/test_gpu_solve.py -m gpu -q -k "multi_gpu or two_ranks or persistent or in_place" > gpurun_out/r2g_pytest.log 2>&1; tail -n 12 gpurun_out/r2g_pytest.log | cut -c1-300
timeout 600 python scripts/spmv_bench.py > gpurun_out/r2g_spmv_1gpu.json 2> gpurun_out/r2g_spmv_1gpu.err; cat gpurun_out/r2g_spmv_1gpu.json | cut -c1-700
timeout 600 $TR --nproc-per-node 2 scripts/spmv_bench.py > gpurun_out/r2g_spmv_2gpu.json 2> gpurun_out/r2g_spmv_2gpu.err; cat gpurun_out/r2g_spmv_2gpu.json | cut -c1-700; tail -n 3 gpurun_out/r2g_spmv_2gpu.err | cut -c1-300
timeout 900 $TR --nproc-per-node 2 bench.py --gpus 2 --steps 5 --warmup 3 --no-aux > gpurun_out/r2g_bench_2gpu.json 2> gpurun_out/r2g_bench_2gpu.err; head -c 600 gpurun_out/r2g_bench_2gpu.json; echo; tail -n 3 gpurun_out/r2g_bench_2gpu.err | cut -c1-300
B200_PROFILE=1 timeout 900 $TR --nproc-per-node 2 bench.py --gpus 2 --steps 2 --warmup 1 --no-aux > gpurun_out/r2g_bench_2gpu_prof.json 2> gpurun_out/r2g_bench_2gpu_prof.err; tail -n 40 gpurun_out/r2g_bench_2gpu_prof.err | cut -c1-200
