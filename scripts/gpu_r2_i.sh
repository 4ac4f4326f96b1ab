#!/bin/bash
# round 2, GPU call I (2 GPUs): native mode with replicated coarse levels; the job that timed out; pattern kernel with register-cached decode
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29513"
timeout 900 python -m pytest tests/test_gpu_solve.py -m gpu -q -k "persistent or in_place or two_ranks" > gpurun_out/r2i_pytest.log 2>&1; tail -n 12 gpurun_out/r2i_pytest.log | cut -c1-400
timeout 900 python scripts/fmt_probe.py --n 256 --levels 0,1 --configs pat8 > gpurun_out/r2i_fmt_probe.jsonl 2> gpurun_out/r2i_fmt_probe.err; tail -n 2 gpurun_out/r2i_fmt_probe.err
B200_PAT_REGS=1 timeout 900 python bench.py --steps 5 --warmup 3 --no-aux --no-cpu-baseline > gpurun_out/r2i_bench_cached.json 2> gpurun_out/r2i_bench_cached.err; head -c 300 gpurun_out/r2i_bench_cached.json; echo; tail -n 2 gpurun_out/r2i_bench_cached.err
timeout 900 python bench.py --steps 5 --warmup 3 --no-aux --no-cpu-baseline > gpurun_out/r2i_bench.json 2> gpurun_out/r2i_bench.err; head -c 300 gpurun_out/r2i_bench.json; echo
HYPRE_B200_VERBOSE=1 timeout 100 hypre_b200/lib/mpirun_shim -np 2 hypre_b200/lib/ij_b200_mp -n 20 20 20 -P 2 1 1 -agg_nl 1 -solver 1 -CF 0 -rlx 0 -w 0.7 -rlx_coarse 0 -ns_coarse 2 > gpurun_out/r2i_job109.log 2>&1; echo "job109 rc=$?"; tail -n 25 gpurun_out/r2i_job109.log | cut -c1-200
timeout 900 $TR --nproc-per-node 2 bench.py --gpus 2 --native --workload lap7_256_cheby --steps 3 --warmup 3 --no-aux > gpurun_out/r2i_bench_c3_native_2gpu.json 2> gpurun_out/r2i_bench_c3_native_2gpu.err; head -c 400 gpurun_out/r2i_bench_c3_native_2gpu.json; echo; tail -n 3 gpurun_out/r2i_bench_c3_native_2gpu.err | cut -c1-300
timeout 900 $TR --nproc-per-node 2 bench.py --gpus 2 --native --weak --workload lap27_128_gs11 --steps 3 --warmup 3 --no-aux > gpurun_out/r2i_bench_c4_native_2gpu.json 2> gpurun_out/r2i_bench_c4_native_2gpu.err; head -c 400 gpurun_out/r2i_bench_c4_native_2gpu.json; echo; tail -n 3 gpurun_out/r2i_bench_c4_native_2gpu.err | cut -c1-300
