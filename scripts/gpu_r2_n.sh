#!/bin/bash
# round 2, GPU call N (2 GPUs): flag-in-data halo exchange / all-reduce, CF-ordered relaxation
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29516"
timeout 900 python -m pytest tests/test_gpu_solve.py -m gpu -q -x -k "multi_gpu or two_ranks or cycle_variants or persistent" > gpurun_out/r2n_pytest.log 2>&1; tail -n 10 gpurun_out/r2n_pytest.log | cut -c1-400
timeout 300 $TR --nproc-per-node 2 scripts/dist_ops.py --dims 256 256 256 --reps 50 > gpurun_out/r2n_ops_2rank.json 2> gpurun_out/r2n_ops_2rank.err; grep "^{" gpurun_out/r2n_ops_2rank.json
timeout 300 $TR --nproc-per-node 2 scripts/dist_ops.py --dims 64 64 64 --reps 200 > gpurun_out/r2n_ops_2rank_64.json 2> gpurun_out/r2n_ops_2rank_64.err; grep "^{" gpurun_out/r2n_ops_2rank_64.json
timeout 300 python scripts/dist_ops.py --dims 64 64 32 --reps 200 > gpurun_out/r2n_ops_1rank_64half.json 2> gpurun_out/r2n_ops_1rank_64half.err; grep "^{" gpurun_out/r2n_ops_1rank_64half.json
timeout 900 $TR --nproc-per-node 2 bench.py --gpus 2 --steps 5 --warmup 3 --no-aux > gpurun_out/r2n_bench_c3_2gpu.json 2> gpurun_out/r2n_bench_c3_2gpu.err; grep "^{" gpurun_out/r2n_bench_c3_2gpu.json | cut -c1-300; tail -n 2 gpurun_out/r2n_bench_c3_2gpu.err | cut -c1-200
timeout 600 $TR --nproc-per-node 2 scripts/spmv_bench.py > gpurun_out/r2n_spmv_c2_2gpu.json 2> gpurun_out/r2n_spmv_c2_2gpu.err; grep "^{" gpurun_out/r2n_spmv_c2_2gpu.json | cut -c1-400
