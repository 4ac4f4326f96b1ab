#!/bin/bash
# round 2, GPU call M (4 GPUs): C3 strong, C2 SpMV, C4 weak (native, hypre's distributed setup) at 4 ranks; the reference's saved jobs (<= 4 ranks) as a gate
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29515"
N=4
free -g | head -2 > gpurun_out/r2m_box.txt; nproc >> gpurun_out/r2m_box.txt; nvidia-smi -L >> gpurun_out/r2m_box.txt
timeout 900 $TR --nproc-per-node $N bench.py --gpus $N --steps 5 --warmup 3 --no-aux > gpurun_out/r2m_bench_c3_${N}gpu.json 2> gpurun_out/r2m_bench_c3_${N}gpu.err; grep "^{" gpurun_out/r2m_bench_c3_${N}gpu.json | cut -c1-250; tail -n 2 gpurun_out/r2m_bench_c3_${N}gpu.err | cut -c1-200
B200_PROFILE=1 timeout 900 $TR --nproc-per-node $N bench.py --gpus $N --steps 2 --warmup 1 --no-aux > gpurun_out/r2m_bench_c3_${N}gpu_prof.json 2> gpurun_out/r2m_bench_c3_${N}gpu_prof.err; grep -A16 "^rank 0" gpurun_out/r2m_bench_c3_${N}gpu_prof.err | cut -c1-150
timeout 600 $TR --nproc-per-node $N scripts/spmv_bench.py > gpurun_out/r2m_spmv_c2_${N}gpu.json 2> gpurun_out/r2m_spmv_c2_${N}gpu.err; grep "^{" gpurun_out/r2m_spmv_c2_${N}gpu.json | cut -c1-500
timeout 900 $TR --nproc-per-node $N bench.py --gpus $N --native --weak --workload lap27_128_gs11 --steps 5 --warmup 3 --no-aux > gpurun_out/r2m_bench_c4_128_native_${N}gpu.json 2> gpurun_out/r2m_bench_c4_128_native_${N}gpu.err; grep "^{" gpurun_out/r2m_bench_c4_128_native_${N}gpu.json | cut -c1-400; tail -n 2 gpurun_out/r2m_bench_c4_128_native_${N}gpu.err | cut -c1-200
timeout 2400 python -m pytest tests/test_gpu_saved_jobs.py -m gpu -q > gpurun_out/r2m_saved_pytest.log 2>&1; tail -n 6 gpurun_out/r2m_saved_pytest.log | cut -c1-400
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/saved_jobs_b200_4gpu.json"))
    print(d["counts"])
    for j in d["jobs"]:
        if j["status"] in ("mismatch", "failed"):
            print(j)
except Exception as e:
    print("no table", e)
PY
