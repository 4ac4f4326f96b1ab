"""Host logic of bench.py that needs no GPU: the workload table stays in step with the golden generator
(tests/golden/gen_bench_golden.py) and with the committed golden histories, and the "parity" block of the
bench line says what it claims (iterations +-1, history within 2e-6 of the printed golden / 1e-9 of the
reference run of the same bench; nothing checked -> ok false)."""
import importlib
import json
import os
import re
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def bench():
    # bench.py fixes OMP_NUM_THREADS for its own process at import: keep this process's value
    saved = os.environ.get("OMP_NUM_THREADS")
    sys.path.insert(0, ROOT)
    try:
        mod = importlib.import_module("bench")
    finally:
        sys.path.remove(ROOT)
        if saved is None:
            os.environ.pop("OMP_NUM_THREADS", None)
        else:
            os.environ["OMP_NUM_THREADS"] = saved
    return mod


@pytest.fixture(scope="module")
def gen():
    sys.path.insert(0, GOLDEN)
    try:
        return importlib.import_module("gen_bench_golden")
    finally:
        sys.path.remove(GOLDEN)


@pytest.fixture(scope="module")
def histories():
    return json.load(open(os.path.join(GOLDEN, "bench_histories.json")))


def test_generator_and_bench_name_the_same_ij_runs(bench, gen):
    for name, args in gen.IJ.items():
        assert name in bench.WORKLOADS, name
        assert bench.WORKLOADS[name][3] == "ij " + args, name
    # every workload the bench can run has a generator line, except the one whose reference solve is too long to
    # regenerate in a bench's time box (BASELINE.md section 5 says so)
    assert set(bench.WORKLOADS) - set(gen.IJ) == {"lap27_320_gs11"}


def test_workload_table_is_consistent_with_its_command_lines(bench):
    flag = {"laplacian": "-laplacian", "laplacian_27pt": "-27pt", "rotate_7pt": "-rotate"}
    for name, (generator, gargs, relax, line, opts) in bench.WORKLOADS.items():
        assert line.startswith("ij " + flag[generator] + " "), name
        dims = [int(v) for v in re.search(r"-n ((?:\d+ ?)+)", line).group(1).split()]
        assert tuple(dims) == tuple(gargs[:len(dims)]), name
        assert "-solver 1" in line and ("-rlx %d" % relax) in line, name
        assert ("-agg_nl 1" in line) == (opts.get("AggNumLevels") == 1), name
        if generator == "rotate_7pt":
            assert "-alpha %g" % gargs[2] in line and "-eps %g" % gargs[3] in line, name


def test_default_workload_is_the_configuration_the_metric_is_quoted_on(bench):
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    text = json.dumps(base)
    # the headline configuration: 7-point 256^3, Chebyshev smoother
    assert "256 256 256" in text and "-rlx 16" in text
    assert bench.WORKLOADS["lap7_256_cheby"][3] == "ij -laplacian -n 256 256 256 -solver 1 -rlx 16"


def test_committed_histories_cover_the_generator_table(gen, histories):
    for name in gen.IJ:
        h = histories[name]
        assert h["iterations"] == len(h["rel_norms"]) >= 5, name
        r = np.asarray(h["rel_norms"])
        assert r[-1] < 1e-8 <= r[-2], name                      # tol 1e-8 reached at the last iteration only
        assert abs(h["final_rel_norm"] - r[-1]) <= 1e-6 * r[-1], name


def test_golden_history_lookup(bench, histories):
    assert bench.golden_history("lap7_256_cheby", None) == histories["lap7_256_cheby"]
    assert bench.golden_history("lap7_256_cheby", 0.5) is None            # scaled-down runs have no golden
    assert bench.golden_history("no_such_workload", None) is None


def _dev(gold, scale=1.0):
    return np.concatenate([[1.0], np.asarray(gold["rel_norms"]) * scale])


def test_parity_block_against_golden_only(bench, histories):
    gold = histories["lap7_100_l1jac"]
    it = gold["iterations"]
    p = bench.parity_block(_dev(gold, 1.0 + 5e-7), it, gold)
    assert p["ok"] is True and p["ref_iterations"] == it and p["tolerance"] == 2e-6
    assert p["max_rel_norm_diff_vs_golden"] == pytest.approx(5e-7, rel=1e-3)
    assert bench.parity_block(_dev(gold, 1.0 + 1e-5), it, gold)["ok"] is False            # history off
    assert bench.parity_block(_dev(gold)[:it - 1], it - 2, gold)["ok"] is False            # two iterations fewer
    assert bench.parity_block(_dev(gold)[:it], it - 1, gold)["ok"] is True                 # +-1 is the contract


def test_parity_block_against_the_reference_run(bench, histories):
    gold = histories["lap7_100_l1jac"]
    it = gold["iterations"]
    full = np.asarray(gold["rel_norms"]) * (1.0 + 3e-7)         # what a full-precision reference run would hold
    ref = {"iterations": it, "rel_norms": full.tolist()}
    p = bench.parity_block(np.concatenate([[1.0], full * (1.0 + 2e-13)]), it, gold, ref)
    assert p["ok"] is True and p["tolerance"] == 1e-9 and p["max_rel_norm_diff"] < 1e-12
    assert "reference" in p
    # within the golden's printed precision but not within 1e-9 of the reference: not ok
    p = bench.parity_block(np.concatenate([[1.0], full * (1.0 + 1e-8)]), it, gold, ref)
    assert p["ok"] is False
    # iteration count differing by more than one
    ref2 = {"iterations": it + 2, "rel_norms": full.tolist() + [1e-9, 1e-10]}
    assert bench.parity_block(np.concatenate([[1.0], full]), it, gold, ref2)["ok"] is False


def test_parity_block_without_anything_to_compare_is_not_ok(bench):
    p = bench.parity_block(np.array([1.0, 0.1, 0.01]), 2, None, None)
    assert p["ok"] is False and p["ref_iterations"] is None and p["max_rel_norm_diff"] is None


def test_committed_ncu_traffic_belongs_to_the_kernel_source_in_the_tree(bench):
    """roofline.traffic is quoted from profiles/r02_ncu_traffic_pat8.json only while spmv.cuh is the file it was
    measured on: the committed pair must match, and a different matrix / format must be refused"""
    t = json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_traffic_pat8.json")))
    assert t["spmv_cuh_sha16"] == bench._sha(os.path.join(ROOT, "hypre_b200", "csrc", "spmv.cuh"))
    m = t["matrix"]
    assert bench.ncu_traffic(m["rows"], m["nnz"], m["value_format"]) == t["spmv_level0"]["traffic_bytes_per_launch"]
    assert bench.ncu_traffic(m["rows"], m["nnz"], 1) is None
    assert bench.ncu_traffic(m["rows"] + 1, m["nnz"], m["value_format"]) is None


def test_launch_summary_is_what_the_script_makes_of_the_committed_launch_list():
    import subprocess
    out = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "summarize_launches.py"),
                          os.path.join(ROOT, "profiles", "r02_launches_lap7_256_cheby.csv")],
                         capture_output=True, text=True, check=True).stdout
    assert out == open(os.path.join(ROOT, "profiles", "r02_launches_lap7_256_cheby_summary.txt")).read()
