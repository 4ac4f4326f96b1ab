#!/usr/bin/env python
"""replay the reference's own regression jobs (src/test/TEST_ij/*.jobs, goldens collected in
tests/golden/ref_saved_goldens.json) on the B200 binding: the unchanged ij.c on N ranks = N GPUs
(hypre_b200/lib/ij_b200_mp over oracle/_ref/mpirun_shim, hypre's own distributed setup, the solve on the
GPUs), compared with the .saved numbers exactly as tests/test_ref_multirank.py compares the reference.

Needs as many GPUs as the widest job has ranks (4): run it as
    gpurun --gpus 4 -- 'python scripts/run_saved_jobs_b200.py > gpurun_out/saved_jobs_b200.json'
Every job ends in one of: pass / mismatch (numbers differ) / rejected (the binding refused an option it has
no device implementation for -- there is no CPU fallback) / failed (non-zero exit, timeout).
Prints one JSON document; exit code 0 when nothing mismatched or failed."""
import argparse
import json
import os
import re
import shlex
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from gen_saved_goldens import parse_output  # noqa: E402

RTOL = 2e-6          # the printed 7 digits (tests/test_ref_multirank.py)


# solver ids the binding runs on the device: 0 AMG, 1/2 PCG, 3/4 GMRES, 16/17 COGMRES, 61/60 FlexGMRES, 51/50 LGMRES,
# 9/10 BiCGSTAB (hypre_glue.c); a job without -solver is ij's default 0
BOUND = {0, 1, 2, 3, 4, 9, 10, 16, 17, 50, 51, 60, 61}


def relevant(job):
    m = re.search(r"-solver (\d+)\b", job["args"])
    return job["suite"] == "smoother" or m is None or int(m.group(1)) in BOUND


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-ranks", type=int, default=4)
    ap.add_argument("--timeout", type=int, default=400)
    ap.add_argument("--all", action="store_true", help="also the jobs of solvers the binding does not interpose (expected: rejected or reference path)")
    ap.add_argument("--match", default="", help="only jobs whose output name contains this")
    a = ap.parse_args()
    jobs = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_saved_goldens.json")))["jobs"]
    launcher = os.path.join(ROOT, "hypre_b200", "lib", "mpirun_shim")
    binary = os.path.join(ROOT, "hypre_b200", "lib", "ij_b200_mp")
    for f in (launcher, binary):
        if not os.path.exists(f):
            sys.exit("missing %s (python -c 'import __graft_entry__ as g; g.build()')" % f)
    env = dict(os.environ, OMP_NUM_THREADS="4", HYPRE_B200_VERBOSE="1")
    report, counts = [], {"pass": 0, "mismatch": 0, "rejected": 0, "failed": 0, "skipped": 0}
    for job in jobs:
        if not (a.all or relevant(job)) or a.match not in job["out"]:
            continue
        entry = {"out": job["out"], "np": job["np"], "args": job["args"]}
        if job["np"] > a.max_ranks:
            entry["status"] = "skipped"
        else:
            try:
                p = subprocess.run([launcher, "-np", str(job["np"]), binary] + shlex.split(job["args"]),
                                   capture_output=True, text=True, timeout=a.timeout, env=env)
                notes = [l for l in (p.stdout + p.stderr).splitlines() if "hypre_b200" in l]
                entry["device_solves"] = sum("solve on the device" in l or "iterations" in l for l in notes)
                got = parse_output(p.stdout.splitlines())
                bad = {}
                for key, want in job["golden"].items():
                    have = got.get(key)
                    ok = have is not None and (have == want if key == "iterations"
                                               else abs(have - want) <= RTOL * abs(want) + 1e-300)
                    if not ok:
                        bad[key] = [have, want]
                rejected = [l for l in notes if any(w in l for w in ("outside the device path", "no device implementation",
                                                                     "must use tol 0", "needs the native", "cannot be combined"))]
                if p.returncode != 0:
                    entry["status"], entry["tail"] = "failed", (p.stderr or p.stdout)[-300:]
                elif rejected:
                    entry["status"], entry["why"] = "rejected", rejected[0][:200]
                elif bad:
                    entry["status"], entry["differs"] = "mismatch", bad
                else:
                    entry["status"] = "pass"
            except subprocess.TimeoutExpired:
                entry["status"], entry["tail"] = "failed", "timeout after %d s" % a.timeout
        counts[entry["status"]] += 1
        report.append(entry)
    print(json.dumps({"counts": counts, "jobs": report}, indent=1))
    return 1 if counts["mismatch"] or counts["failed"] else 0


if __name__ == "__main__":
    sys.exit(main())
