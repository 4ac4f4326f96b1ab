#!/bin/bash
# round 2, one GPU: the GPU test suite, then every BASELINE configuration that fits one GPU (C1, C2, C3, C4 at 128^3
# and 320^3, C5), the per-kernel breakdown of C3 / C1 (B200_PROFILE=1) and the storage-format table per level.
#   gpurun --timeout 3000 -- 'bash scripts/gpu_r2_single.sh'      -> profiles/r02_bench_c*.json, r02_spmv_c2_1gpu.json,
#   r02_kernel_times_c{1,3}.txt, r02_fmt_probe.jsonl, r02_pytest.log (copied from gpurun_out/ by hand)
mkdir -p gpurun_out; O=gpurun_out/r02
free -g | head -2 > ${O}_box.txt; nproc >> ${O}_box.txt; nvidia-smi -L >> ${O}_box.txt
timeout 1200 python -m pytest tests -m gpu -q > ${O}_pytest.log 2>&1; tail -n 8 ${O}_pytest.log | cut -c1-300
timeout 900 python bench.py --steps 10 --warmup 3 > ${O}_bench_c3.json 2> ${O}_bench_c3.err; head -c 300 ${O}_bench_c3.json; echo
B200_PROFILE=1 timeout 600 python bench.py --steps 2 --warmup 1 --no-aux --no-cpu-baseline > /dev/null 2> ${O}_kernel_times_c3.txt
timeout 600 python bench.py --workload lap7_100_l1jac --steps 20 --warmup 3 --no-aux > ${O}_bench_c1.json 2> ${O}_bench_c1.err; head -c 300 ${O}_bench_c1.json; echo
B200_PROFILE=1 timeout 600 python bench.py --workload lap7_100_l1jac --steps 2 --warmup 1 --no-aux --no-cpu-baseline > /dev/null 2> ${O}_kernel_times_c1.txt
timeout 900 python scripts/fmt_probe.py --n 256 --levels 0,1,2 --configs pat8,val_zpad,val_markers,plain > ${O}_fmt_probe.jsonl 2> ${O}_fmt_probe.err
timeout 600 python scripts/spmv_bench.py > ${O}_spmv_c2_1gpu.json 2> ${O}_spmv_c2_1gpu.err; cut -c1-500 ${O}_spmv_c2_1gpu.json
timeout 900 python bench.py --workload lap27_128_gs11 --steps 10 --warmup 3 --no-aux > ${O}_bench_c4_128.json 2> ${O}_bench_c4_128.err; head -c 300 ${O}_bench_c4_128.json; echo
timeout 1500 python bench.py --workload rot_8192_l1jac_agg1 --steps 5 --warmup 3 --no-aux > ${O}_bench_c5.json 2> ${O}_bench_c5.err; head -c 300 ${O}_bench_c5.json; echo
timeout 1500 python bench.py --workload lap27_320_gs11 --steps 5 --warmup 3 --no-aux --no-cpu-baseline > ${O}_bench_c4_320.json 2> ${O}_bench_c4_320.err; head -c 300 ${O}_bench_c4_320.json; echo
