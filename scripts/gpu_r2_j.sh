#!/bin/bash
# round 2, GPU call J (1 GPU): full GPU test suite, the five BASELINE configs that fit one GPU
mkdir -p gpurun_out
free -g | head -2 > gpurun_out/r2j_box.txt; nproc >> gpurun_out/r2j_box.txt; nvidia-smi -L >> gpurun_out/r2j_box.txt
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r2j_pytest.log 2>&1; tail -n 8 gpurun_out/r2j_pytest.log | cut -c1-300
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2j_bench_c3.json 2> gpurun_out/r2j_bench_c3.err; tail -n 2 gpurun_out/r2j_bench_c3.err; head -c 300 gpurun_out/r2j_bench_c3.json; echo
B200_PROFILE=1 timeout 600 python bench.py --steps 2 --warmup 1 --no-aux --no-cpu-baseline > gpurun_out/r2j_bench_c3_prof.json 2> gpurun_out/r2j_bench_c3_prof.err; tail -n 14 gpurun_out/r2j_bench_c3_prof.err
timeout 600 python bench.py --workload lap7_100_l1jac --steps 20 --warmup 3 --no-aux > gpurun_out/r2j_bench_c1.json 2> gpurun_out/r2j_bench_c1.err; tail -n 2 gpurun_out/r2j_bench_c1.err; head -c 300 gpurun_out/r2j_bench_c1.json; echo
B200_PROFILE=1 timeout 600 python bench.py --workload lap7_100_l1jac --steps 2 --warmup 1 --no-aux --no-cpu-baseline > gpurun_out/r2j_bench_c1_prof.json 2> gpurun_out/r2j_bench_c1_prof.err; tail -n 16 gpurun_out/r2j_bench_c1_prof.err
timeout 900 python scripts/fmt_probe.py --n 256 --levels 0,1,2 --configs pat8,val_zpad,val_markers,plain > gpurun_out/r2j_fmt_probe.jsonl 2> gpurun_out/r2j_fmt_probe.err; tail -n 2 gpurun_out/r2j_fmt_probe.err
timeout 600 python scripts/spmv_bench.py > gpurun_out/r2j_spmv_c2_1gpu.json 2> gpurun_out/r2j_spmv_c2_1gpu.err; cut -c1-500 gpurun_out/r2j_spmv_c2_1gpu.json
timeout 900 python bench.py --workload lap27_128_gs11 --steps 10 --warmup 3 --no-aux > gpurun_out/r2j_bench_c4_128.json 2> gpurun_out/r2j_bench_c4_128.err; tail -n 2 gpurun_out/r2j_bench_c4_128.err; head -c 300 gpurun_out/r2j_bench_c4_128.json; echo
timeout 1500 python bench.py --workload rot_8192_l1jac_agg1 --steps 5 --warmup 3 --no-aux > gpurun_out/r2j_bench_c5.json 2> gpurun_out/r2j_bench_c5.err; tail -n 3 gpurun_out/r2j_bench_c5.err | cut -c1-300; head -c 400 gpurun_out/r2j_bench_c5.json; echo
timeout 1500 python bench.py --workload lap27_320_gs11 --steps 5 --warmup 3 --no-aux --no-cpu-baseline > gpurun_out/r2j_bench_c4_320.json 2> gpurun_out/r2j_bench_c4_320.err; tail -n 3 gpurun_out/r2j_bench_c4_320.err | cut -c1-300; head -c 400 gpurun_out/r2j_bench_c4_320.json; echo
