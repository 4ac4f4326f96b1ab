#!/bin/bash
# compressed sliced-ELL storage: parity tests, operator table per format / pipelining variant, bench A/B
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_c.log 2>&1; tail -n 5 gpurun_out/pytest_gpu_c.log
for cfg in "-1 0" "-1 1" "-1 2" "-1 3" "3 2" "3 3"; do
  set -- $cfg
  python scripts/prof_ops.py --compress $1 --pipe $2 > gpurun_out/ops_lap7_256_c$1_p$2.json 2> gpurun_out/ops_c.err || tail -n 3 gpurun_out/ops_c.err
  cut -c 160-1000 gpurun_out/ops_lap7_256_c$1_p$2.json
done
for cfg in "-1 0" "-1 2" "-1 3"; do
  set -- $cfg
  python scripts/prof_ops.py --gen laplacian_27pt --n 200 --compress $1 --pipe $2 > gpurun_out/ops_lap27_200_c$1_p$2.json 2> gpurun_out/ops_c.err || tail -n 3 gpurun_out/ops_c.err
  cut -c 160-1000 gpurun_out/ops_lap27_200_c$1_p$2.json
done
for p in 2 3; do
B200_SELLC_PIPE=$p python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench1_256_c_p$p.log 2>&1; tail -n 1 gpurun_out/bench1_256_c_p$p.log | cut -c 1-2500
done
