#!/bin/bash
# round 2, one GPU: validation of the committed tree -- what the driver runs at round end.
#   the allocation-free entry point alone in a fresh process, the whole GPU suite (-x), smoke(), both bench arms
#   -> profiles/r02_noalloc_alone.log, r02_pytest_gpu_final.log, r02_smoke_final.log, r02_bench_final_{default,reference}.json
mkdir -p gpurun_out; O=gpurun_out/r02
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > ${O}_noalloc_alone.log 2>&1; tail -n 3 ${O}_noalloc_alone.log | cut -c1-300
timeout 1200 python -m pytest tests -m gpu -x -q > ${O}_pytest_gpu_final.log 2>&1; tail -n 6 ${O}_pytest_gpu_final.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" > ${O}_smoke_final.log 2>&1; tail -n 2 ${O}_smoke_final.log
timeout 900 python bench.py > ${O}_bench_final_default.json 2> ${O}_bench_final_default.err; head -c 300 ${O}_bench_final_default.json; echo
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > ${O}_bench_final_reference.json 2> ${O}_bench_final_reference.err; head -c 300 ${O}_bench_final_reference.json; echo
