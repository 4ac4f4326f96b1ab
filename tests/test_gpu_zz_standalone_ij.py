"""GPU test, collected last: the reference's unchanged ij.c as a stand-alone BoomerAMG solver (`ij -solver 0`,
the form of most TEST_ij/smoother.jobs lines) on the binding.  hypre_glue.c interposes hypre_BoomerAMGSolve
(src/parcsr_ls/par_amg_solve.c:21-424): cycles on the device (b200_amg_solve), table and closing statistics
printed by the binding.  Compared with the same driver on the reference library: iteration count, final
relative residual, average convergence factor and the three complexities ij prints."""
import os
import re
import subprocess

import pytest

from hypre_b200 import IJ_B200, IJ_REF


pytestmark = [pytest.mark.gpu]

KEYS = {"iterations": r"BoomerAMG Iterations = (\d+)",
        "final_rel": r"Final Relative Residual Norm = (\S+)",
        "avg_conv": r"Average Convergence Factor = (\S+)",
        "grid": r"Complexity:\s+grid = (\S+)",
        "operator": r"operator = (\S+)",
        "cycle": r"cycle = (\S+)"}


def run(binary, args):
    env = dict(os.environ, HYPRE_B200_VERBOSE="1")
    r = subprocess.run([binary] + args.split(), capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    got = {}
    for k, pat in KEYS.items():
        m = re.findall(pat, r.stdout)          # the last match: the setup prints complexities of its own
        assert m, (k, r.stdout[-1500:])
        got[k] = float(m[-1])
    rows = re.findall(r"^\s+Cycle\s+\d+\s+(\S+e[-+]\d+)\s+\S+\s+(\S+e[-+]\d+)", r.stdout, re.M)
    return got, [(float(a), float(b)) for a, b in rows], r.stdout + r.stderr


@pytest.mark.parametrize("args", ["-laplacian -n 24 24 24 -solver 0 -rlx 18",
                                  "-laplacian -n 20 20 20 -solver 0 -rlx 16",
                                  "-27pt -n 16 16 16 -solver 0 -rlx 11",
                                  "-laplacian -n 20 20 20 -solver 0 -rlx 18 -mu 2"])
def test_unchanged_ij_standalone_amg_on_the_device(args):
    if not (os.path.exists(IJ_B200) and os.path.exists(IJ_REF)):
        pytest.skip("ij binaries not built")
    dev, rows_d, out_d = run(IJ_B200, args)
    ref, rows_r, _ = run(IJ_REF, args)
    assert "BoomerAMG solve on the device" in out_d            # not the reference's CPU loop
    assert dev["iterations"] == ref["iterations"]
    assert len(rows_d) == len(rows_r) == int(ref["iterations"])
    for (rd, reld), (rr, relr) in zip(rows_d, rows_r):
        # the residual is recomputed from x: near convergence nine digits cancel (see test_gpu_solve.py)
        assert abs(rd - rr) <= 2e-6 * rr + 1e-14 and abs(reld - relr) <= 2e-6 * relr + 1e-15
    assert abs(dev["final_rel"] - ref["final_rel"]) <= 2e-6 * ref["final_rel"] + 1e-15
    for k in ("avg_conv", "grid", "operator", "cycle"):
        assert abs(dev[k] - ref[k]) <= 2e-6 * abs(ref[k]) + 1e-6, (k, dev[k], ref[k])
