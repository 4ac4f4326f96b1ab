#!/bin/bash
# round 2, GPU call F: persistent coarse-level kernel (tail.cuh)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2f_pytest.log 2>&1; tail -n 25 gpurun_out/r2f_pytest.log | cut -c1-300
timeout 900 python bench.py --steps 5 --warmup 3 --no-aux --no-cpu-baseline > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err; tail -n 3 gpurun_out/r2f_bench.err; head -c 300 gpurun_out/r2f_bench.json; echo
B200_TAIL_NNZ=0 timeout 900 python bench.py --steps 5 --warmup 3 --no-aux --no-cpu-baseline > gpurun_out/r2f_bench_notail.json 2> gpurun_out/r2f_bench_notail.err; head -c 300 gpurun_out/r2f_bench_notail.json; echo
timeout 900 python bench.py --workload lap7_100_l1jac --steps 20 --warmup 3 --no-aux --no-cpu-baseline > gpurun_out/r2f_bench_c1.json 2> gpurun_out/r2f_bench_c1.err; head -c 300 gpurun_out/r2f_bench_c1.json; echo
B200_TAIL_NNZ=0 timeout 900 python bench.py --workload lap7_100_l1jac --steps 20 --warmup 3 --no-aux --no-cpu-baseline > gpurun_out/r2f_bench_c1_notail.json 2> gpurun_out/r2f_bench_c1_notail.err; head -c 300 gpurun_out/r2f_bench_c1_notail.json; echo
B200_PROFILE=1 timeout 600 python bench.py --steps 2 --warmup 1 --no-aux --no-cpu-baseline > gpurun_out/r2f_bench_prof.json 2> gpurun_out/r2f_bench_prof.err; tail -n 16 gpurun_out/r2f_bench_prof.err
