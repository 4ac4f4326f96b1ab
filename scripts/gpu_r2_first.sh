#!/bin/bash
# first GPU call of round 2: everything written without GPU minutes at the end of round 1 gets its first run
#   1. the whole GPU suite (the new files are collected last: test_gpu_zz_krylov.py, test_gpu_zz_standalone_ij.py)
#   2. candidate kernels of the compressed row loop (variants 4, 5), opt-in
#   3. smoke + the contract bench line + GMRES / BiCGSTAB timings
#   gpurun --timeout 1500 -- 'bash scripts/gpu_r2_first.sh'
mkdir -p gpurun_out
B200_TEST_UNVERIFIED=1 timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2_pytest_gpu.log 2>&1; tail -n 15 gpurun_out/r2_pytest_gpu.log
B200_TEST_CANDIDATES=1 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "compressed" > gpurun_out/r2_pytest_candidates.log 2>&1; tail -n 5 gpurun_out/r2_pytest_candidates.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; tail -n 3 gpurun_out/r2_smoke.log
python bench.py > gpurun_out/r2_bench1_256.log 2>&1; tail -n 1 gpurun_out/r2_bench1_256.log | cut -c 1-700
python scripts/bench_krylov.py --n 128 > gpurun_out/r2_bench_krylov_128.jsonl 2>&1; cat gpurun_out/r2_bench_krylov_128.jsonl | cut -c 1-400
python scripts/bench_krylov.py --n 256 --steps 2 > gpurun_out/r2_bench_krylov_256.jsonl 2>&1; cat gpurun_out/r2_bench_krylov_256.jsonl | cut -c 1-400
B200_SELLC_PIPE=5 python bench.py --no-cpu-baseline --no-aux > gpurun_out/r2_bench1_256_pipe5.log 2>&1; tail -n 1 gpurun_out/r2_bench1_256_pipe5.log | cut -c 1-400
# launch list of the Krylov solvers (shares of the Gram-Schmidt / update kernels per solve), only after the same
# command has exited 0 without ncu
python scripts/bench_krylov.py --n 128 --steps 1 --warmup 1 > gpurun_out/r2_krylov_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r2_launches_krylov.csv python scripts/bench_krylov.py --n 128 --steps 1 --warmup 1 > gpurun_out/r2_ncu_krylov.log 2>&1
tail -n 1 gpurun_out/r2_ncu_krylov.log | cut -c 1-200
# with 4 GPUs (gpurun --gpus 4): the reference's own regression jobs on the binding, N ranks = N GPUs
#   python scripts/run_saved_jobs_b200.py > gpurun_out/r2_saved_jobs_b200.json
