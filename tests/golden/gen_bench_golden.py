#!/usr/bin/env python
"""Per-iteration residual histories of the REFERENCE for bench.py's workloads at FULL size
(tests/golden/bench_histories.json): one run of oracle/_ref/ij_ref (unmodified src/test/ij.c over the
unmodified reference library) per workload, table of hypre_PCGSolve parsed (src/krylov/pcg.c:772-792).
bench.py compares the device solve's history with these at every N (its "parity" block); nothing reads
/root/reference at bench time.

    python tests/golden/gen_bench_golden.py lap7_256_cheby [more workloads ...]
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

from gen_golden import run  # noqa: E402

# bench.py WORKLOADS -> ij arguments (kept in step with bench.py by tests/test_bench_host.py)
IJ = {
    "lap7_256_cheby": "-laplacian -n 256 256 256 -solver 1 -rlx 16",
    "lap7_256_l1jac": "-laplacian -n 256 256 256 -solver 1 -rlx 18",
    "lap7_100_l1jac": "-laplacian -n 100 100 100 -solver 1 -rlx 18",
    "lap27_128_gs11": "-27pt -n 128 128 128 -solver 1 -rlx 11",
    "lap27_200_gs11": "-27pt -n 200 200 200 -solver 1 -rlx 11",
    "lap7_64_cheby": "-laplacian -n 64 64 64 -solver 1 -rlx 16",
    "rot_2048_l1jac_agg1": "-rotate -n 2048 2048 -alpha 45 -eps 0.001 -solver 1 -rlx 18 -agg_nl 1",
    "rot_8192_l1jac_agg1": "-rotate -n 8192 8192 -alpha 45 -eps 0.001 -solver 1 -rlx 18 -agg_nl 1",
}


def main():
    path = os.path.join(HERE, "bench_histories.json")
    gold = json.load(open(path)) if os.path.exists(path) else {}
    for name in sys.argv[1:] or ["lap7_256_cheby"]:
        gold[name] = run(IJ[name])
        print(name, gold[name]["iterations"], gold[name]["final_rel_norm"], flush=True)
        gold["_meta"] = {"generator": "tests/golden/gen_bench_golden.py",
                         "reference": "hypre 2.32.0 (PaulMullowney/hypre), host build of oracle/Makefile, ij_ref",
                         "printed_precision": "7 significant digits (%e)"}
        json.dump(gold, open(path, "w"), indent=1)


if __name__ == "__main__":
    main()
