#!/bin/bash
# round 2, GPU call H (2 GPUs): fixed tests, the reference's saved jobs with <= 2 ranks as a gate, native weak-scaling path (C4) at 2 ranks
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29512"
timeout 900 python -m pytest tests/test_gpu_solve.py -m gpu -q -k "persistent or in_place" > gpurun_out/r2h_pytest.log 2>&1; tail -n 12 gpurun_out/r2h_pytest.log | cut -c1-400
timeout 1500 python -m pytest tests/test_gpu_saved_jobs.py -m gpu -q > gpurun_out/r2h_saved_pytest.log 2>&1; tail -n 6 gpurun_out/r2h_saved_pytest.log | cut -c1-400
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/saved_jobs_b200_2gpu.json"))
    print(d["counts"])
    for j in d["jobs"]:
        if j["status"] in ("mismatch", "failed"):
            print(j)
except Exception as e:
    print("no table", e)
PY
timeout 900 $TR --nproc-per-node 2 bench.py --gpus 2 --native --weak --workload lap27_128_gs11 --steps 3 --warmup 3 --no-aux > gpurun_out/r2h_bench_c4_native_2gpu.json 2> gpurun_out/r2h_bench_c4_native_2gpu.err; head -c 900 gpurun_out/r2h_bench_c4_native_2gpu.json; echo; tail -n 5 gpurun_out/r2h_bench_c4_native_2gpu.err | cut -c1-300
timeout 900 $TR --nproc-per-node 2 bench.py --gpus 2 --native --workload lap7_256_cheby --steps 3 --warmup 3 --no-aux > gpurun_out/r2h_bench_c3_native_2gpu.json 2> gpurun_out/r2h_bench_c3_native_2gpu.err; head -c 900 gpurun_out/r2h_bench_c3_native_2gpu.json; echo; tail -n 5 gpurun_out/r2h_bench_c3_native_2gpu.err | cut -c1-300
