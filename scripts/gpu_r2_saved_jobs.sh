#!/bin/bash
# round 2, N GPUs (default 4): the reference's regression jobs (src/test/TEST_ij/{solvers,smoother}.jobs, recorded with
# mpirun -np 2..4) through the unchanged ij on the device backend; iteration counts must equal the .saved files.
# Twelve minutes at 4 GPUs: charged N x, run it once per tree.
#   gpurun --gpus 4 --timeout 2400 -- 'bash scripts/gpu_r2_saved_jobs.sh 4'   -> profiles/r02_saved_jobs_b200_4gpu.json
N=${1:-4}
mkdir -p gpurun_out
timeout 2200 python scripts/run_saved_jobs_b200.py --max-ranks $N > gpurun_out/r02_saved_jobs_b200_${N}gpu.json 2> gpurun_out/r02_saved_jobs.err
python - $N <<'PY'
import json, sys
d = json.load(open("gpurun_out/r02_saved_jobs_b200_%sgpu.json" % sys.argv[1]))
print(d["counts"])
for j in d["jobs"]:
    if j["status"] in ("mismatch", "failed"):
        print(j)
PY
