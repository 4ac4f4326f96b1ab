#!/bin/bash
# round 2, GPU call R (4 GPUs): the reference's smoother.jobs (3-4 ranks) through the device backend after the CF-ordering / gathered-level fixes
mkdir -p gpurun_out
timeout 1200 python scripts/run_saved_jobs_b200.py --max-ranks 4 --match smoother > gpurun_out/r2r_saved_smoother_4gpu.json 2> gpurun_out/r2r_saved_smoother.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2r_saved_smoother_4gpu.json"))
print(d["counts"])
for j in d["jobs"]:
    if j["status"] in ("mismatch", "failed"):
        print(j)
    elif j["status"] == "pass":
        print("pass", j["out"], j["args"][:90])
PY
