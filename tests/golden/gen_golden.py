#!/usr/bin/env python
"""Generates tests/golden/pcg_histories.json by running the REFERENCE's own driver
(oracle/_ref/ij_ref = unmodified src/test/ij.c linked against the unmodified reference
library) and parsing the per-iteration table hypre_PCGSolve prints at print_level > 1
(src/krylov/pcg.c:772-792).  Run in the build container (needs oracle/_ref built):

    python tests/golden/gen_golden.py

The committed JSON is what the tests read; nothing reads /root/reference at test time.
"""
import json
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)

from oracle_util import IJ_ARGS  # noqa: E402

IJ_REF = os.path.join(ROOT, "oracle", "_ref", "ij_ref")


def run(args):
    env = dict(os.environ, OMP_NUM_THREADS=os.environ.get("OMP_NUM_THREADS", "8"))
    out = subprocess.run([IJ_REF] + args.split(), capture_output=True, text=True, env=env, check=True).stdout
    rel, norms = [], []
    for m in re.finditer(r"^\s+(\d+)\s+([0-9.e+-]+)\s+([0-9.]+)\s+([0-9.e+-]+)\s*$", out, re.M):
        norms.append(float(m.group(2)))
        rel.append(float(m.group(4)))
    its = int(re.search(r"Iterations = (\d+)", out).group(1))
    fin = float(re.search(r"Final Relative Residual Norm = ([0-9.e+-]+)", out).group(1))
    lev = re.search(r"Num levels = (\d+)", out)
    cx = re.search(r"operator = ([0-9.]+)", out)
    assert len(rel) == its, (args, len(rel), its)
    return {"ij_args": args, "iterations": its, "final_rel_norm": fin, "rel_norms": rel, "norms": norms,
            "num_levels": int(lev.group(1)) if lev else None,
            "operator_complexity": float(cx.group(1)) if cx else None}


def main():
    gold = {name: run(args) for name, args in IJ_ARGS.items()}
    gold["_meta"] = {"generator": "tests/golden/gen_golden.py", "reference": "hypre 2.32.0 (PaulMullowney/hypre), "
                     "host build: HYPRE_SEQUENTIAL + OpenMP, oracle/Makefile",
                     "printed_precision": "7 significant digits (%e)"}
    with open(os.path.join(HERE, "pcg_histories.json"), "w") as f:
        json.dump(gold, f, indent=1)
    for k, v in gold.items():
        if k != "_meta":
            print(k, v["iterations"], v["final_rel_norm"])


if __name__ == "__main__":
    main()
