#!/usr/bin/env python
"""Collect the reference's own multi-rank golden results for the ij driver into
tests/golden/ref_saved_goldens.json.

Source: /root/reference/src/test/TEST_ij/<suite>.jobs (the command lines, `mpirun -np N ./ij ...`)
and <suite>.saved (what the reference's regression harness src/test/runtest.sh compares against:
iteration counts, final relative residual norms, average convergence factors, complexities).
Jobs that read matrix files shipped with the reference's test data are left out (the files do not
travel).  tests/test_ref_multirank.py replays the jobs on N ranks with the reference compiled by
oracle/Makefile against hypre_b200/host/mpishim and checks every number.

    python tests/golden/gen_saved_goldens.py [suite ...]
"""
import json
import os
import re
import sys

REF = "/root/reference/src/test/TEST_ij"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_saved_goldens.json")
SUITES = ["smoother", "solvers", "default", "coarsening", "interp"]

PATTERNS = {
    "iterations": re.compile(r"^\s*(?:[A-Za-z-]+ )?Iterations\s*=\s*(\d+)"),
    "final_rel_residual": re.compile(r"^\s*Final Relative Residual Norm\s*=\s*(\S+)"),
    "avg_conv_factor": re.compile(r"^\s*Average Convergence Factor\s*=\s*(\S+)"),
    "complexity_grid": re.compile(r"^\s*Complexity:\s*grid\s*=\s*(\S+)"),
    "complexity_operator": re.compile(r"^\s*operator\s*=\s*(\S+)"),
    "complexity_cycle": re.compile(r"^\s*cycle\s*=\s*(\S+)"),
}


def parse_output(lines):
    """the LAST occurrence of every quantity (ij prints setup statistics first, the solve summary last)"""
    res = {}
    for line in lines:
        for key, pat in PATTERNS.items():
            m = pat.match(line)
            if m:
                res[key] = int(m.group(1)) if key == "iterations" else float(m.group(1))
    return res


def parse_jobs(path):
    txt = open(path).read().replace("\\\n", " ")
    jobs = []
    for line in txt.splitlines():
        m = re.match(r"\s*mpirun\s+-np\s+(\d+)\s+\./ij\s*(.*?)\s*>\s*(\S+)\s*$", line)
        if m:
            jobs.append({"np": int(m.group(1)), "args": " ".join(m.group(2).split()), "out": m.group(3)})
    return jobs


def parse_saved(path):
    blocks, cur = {}, None
    for line in open(path):
        m = re.match(r"# Output file: (\S+)", line)
        if m:
            cur = m.group(1)
            blocks[cur] = []
        elif cur is not None:
            blocks[cur].append(line)
    return {k: parse_output(v) for k, v in blocks.items()}


def main():
    suites = sys.argv[1:] or SUITES
    entries = []
    for s in suites:
        jobs_p, saved_p = os.path.join(REF, s + ".jobs"), os.path.join(REF, s + ".saved")
        if not (os.path.exists(jobs_p) and os.path.exists(saved_p)):
            continue
        saved = parse_saved(saved_p)
        for j in parse_jobs(jobs_p):
            if re.search(r"-fromfile|-fromonecsrfile|-rhsfromfile|-rhsparcsrfile|-x0fromfile|-auxfromfile", j["args"]):
                continue
            g = saved.get(j["out"])
            if not g:
                continue
            entries.append({"suite": s, "source": "src/test/TEST_ij/%s.saved" % s, **j, "golden": g})
    json.dump({"reference": "hypre 2.32.0, src/test/TEST_ij (runtest.sh goldens)", "jobs": entries}, open(OUT, "w"), indent=1)
    print("wrote %d jobs to %s" % (len(entries), OUT))


if __name__ == "__main__":
    main()
