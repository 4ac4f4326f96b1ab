#!/bin/bash
# round 2, GPU call P (2 GPUs): small distributed cases with empty ranks on coarse levels; wide-gather variant of the pattern kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_solve.py -m gpu -q -k "two_ranks" > gpurun_out/r2p_pytest.log 2>&1; tail -n 10 gpurun_out/r2p_pytest.log | cut -c1-400
timeout 900 python scripts/fmt_probe.py --n 256 --levels 0,1 --configs pat8 > gpurun_out/r2p_fmt_probe.jsonl 2> gpurun_out/r2p_fmt_probe.err; tail -n 2 gpurun_out/r2p_fmt_probe.err
python - <<'PY'
import json
for l in open("gpurun_out/r2p_fmt_probe.jsonl"):
    d = json.loads(l)
    print("L%d %s wide=%d | " % (d["level"], d["config"], d["pat_regs"] & 1) + " | ".join("%s %6.1fus" % (k, d[k]["us"]) for k in ("spmv", "residual", "l1jacobi", "cheby2")), [v for k, v in d.items() if k.endswith("_identical")])
PY
B200_PAT_REGS=1 timeout 900 python bench.py --steps 5 --warmup 3 --no-aux --no-cpu-baseline > gpurun_out/r2p_bench_wide.json 2> gpurun_out/r2p_bench_wide.err; head -c 300 gpurun_out/r2p_bench_wide.json; echo
