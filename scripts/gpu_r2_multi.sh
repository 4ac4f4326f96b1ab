#!/bin/bash
# round 2, N GPUs of one box (N = 2, 4, 8): multi-rank parity tests, C3 strong scaling, C2 SpMV, C4 weak scaling in
# native mode (hypre's distributed setup on the N ranks, 128^3 per GPU), level-0 operator table at 2 ranks.
#   gpurun --gpus N --timeout 2400 -- 'bash scripts/gpu_r2_multi.sh N'
#   -> profiles/r02_bench_c3_{N}gpu.json, r02_spmv_c2_{N}gpu.json, r02_bench_c4_128_native_{N}gpu.json,
#      r02_kernel_times_c3_{N}gpu.txt, r02_ops_2rank*.json, r02_pytest_2gpu*.log
N=${1:-2}
mkdir -p gpurun_out; O=gpurun_out/r02
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29511 --nproc-per-node $N"
free -g | head -2 > ${O}_box_${N}gpu.txt; nproc >> ${O}_box_${N}gpu.txt; nvidia-smi -L >> ${O}_box_${N}gpu.txt
if [ "$N" = 2 ]; then
timeout 900 python -m pytest tests/test_gpu_solve.py -m gpu -q -k "multi_gpu or two_ranks or cycle_variants or persistent or in_place" > ${O}_pytest_2gpu.log 2>&1; tail -n 10 ${O}_pytest_2gpu.log | cut -c1-300
timeout 300 $TR scripts/dist_ops.py --dims 256 256 256 --reps 50 > ${O}_ops_2rank.json 2> ${O}_ops_2rank.err; grep "^{" ${O}_ops_2rank.json
timeout 300 python scripts/dist_ops.py --dims 256 256 128 --reps 50 > ${O}_ops_1rank_halfrows.json 2> /dev/null; cat ${O}_ops_1rank_halfrows.json
B200_HALO=nccl timeout 300 $TR scripts/dist_ops.py --dims 256 256 256 --reps 50 > ${O}_ops_2rank_nccl.json 2> /dev/null; grep "^{" ${O}_ops_2rank_nccl.json
fi
timeout 900 $TR bench.py --gpus $N --steps 5 --warmup 3 --no-aux > ${O}_bench_c3_${N}gpu.json 2> ${O}_bench_c3_${N}gpu.err; grep "^{" ${O}_bench_c3_${N}gpu.json | cut -c1-260
B200_PROFILE=1 timeout 900 $TR bench.py --gpus $N --steps 2 --warmup 1 --no-aux > /dev/null 2> ${O}_kernel_times_c3_${N}gpu.txt
timeout 600 $TR scripts/spmv_bench.py > ${O}_spmv_c2_${N}gpu.json 2> ${O}_spmv_c2_${N}gpu.err; grep "^{" ${O}_spmv_c2_${N}gpu.json | cut -c1-400
timeout 900 $TR bench.py --gpus $N --native --weak --workload lap27_128_gs11 --steps 5 --warmup 3 --no-aux > ${O}_bench_c4_128_native_${N}gpu.json 2> ${O}_bench_c4_128_native_${N}gpu.err; grep "^{" ${O}_bench_c4_128_native_${N}gpu.json | cut -c1-400
