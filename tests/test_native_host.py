"""CPU tests of the native multi-rank mode as bench.py uses it: ranks that were NOT started by mpirun_shim
(torchrun starts them) meet through a directory (SHIM_DIR, hypre_b200/host/mpishim/mpishim.c: rendezvous) and
hypre_b200/hypre.py drives the MPI build of the reference on them.  With the device path off the result must
be what the reference's own ij driver prints on the same process grid."""
import os
import re
import subprocess
import sys
import tempfile
import time

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref")
HAVE = all(os.path.exists(os.path.join(BIN, f)) for f in ("mpirun_shim", "ij_ref_mp", "shim_selftest")) and \
    os.path.exists(os.path.join(ROOT, "hypre_b200", "lib", "libHYPRE_b200_mp.so"))
pytestmark = pytest.mark.skipif(not HAVE, reason="multi-rank build absent")


def launch(n, argv, stagger=0.0, extra_env=None, timeout=300):
    """n processes with torchrun's environment (RANK, WORLD_SIZE) and a fresh SHIM_DIR; rank 0 may start last"""
    with tempfile.TemporaryDirectory(prefix="shimrdv") as d:
        procs = []
        for r in reversed(range(n)):
            env = dict(os.environ, SHIM_DIR=d, RANK=str(r), WORLD_SIZE=str(n), LOCAL_RANK=str(r), OMP_NUM_THREADS="1")
            env.pop("SHIM_FDS", None); env.pop("SHIM_RANK", None); env.pop("SHIM_SIZE", None)
            env.update(extra_env or {})
            procs.append((r, subprocess.Popen(argv, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=env)))
            if stagger:
                time.sleep(stagger)
        out = {}
        for r, p in procs:
            so, se = p.communicate(timeout=timeout)
            out[r] = (p.returncode, so, se)
        return out


@pytest.mark.parametrize("n", [2, 5])
def test_directory_rendezvous_runs_the_selftest(n):
    out = launch(n, [os.path.join(BIN, "shim_selftest")], stagger=0.15 if n == 2 else 0.0, timeout=120)
    assert all(rc == 0 for rc, _, _ in out.values()), {r: v[2][-300:] for r, v in out.items()}
    assert "SHIM_OK %d ranks" % n in out[0][1]


@pytest.mark.parametrize("grid,dims,rlx,stencil", [((2, 1, 1), (20, 20, 20), 18, "7"),
                                                   ((2, 2, 1), (18, 16, 12), 16, "7"),
                                                   ((1, 3, 1), (10, 15, 10), 18, "27")])
def test_python_native_mode_equals_the_ij_driver(grid, dims, rlx, stencil):
    n = grid[0] * grid[1] * grid[2]
    argv = [sys.executable, os.path.join(ROOT, "tests", "native_worker.py")] + [str(v) for v in grid + dims] + [str(rlx)]
    if stencil == "27":
        argv.append("27")
    out = launch(n, argv)
    assert all(rc == 0 for rc, _, _ in out.values()), {r: v[2][-400:] for r, v in out.items()}
    m = re.search(r"NATIVE_OK ranks=(\d+) rows=(\d+) levels=(\d+) iterations=(\d+) final=(\S+)", out[0][1])
    assert m and int(m.group(1)) == n and int(m.group(2)) == dims[0] * dims[1] * dims[2]
    ij = subprocess.run([os.path.join(BIN, "mpirun_shim"), "-np", str(n), os.path.join(BIN, "ij_ref_mp"),
                         "-27pt" if stencil == "27" else "-laplacian", "-n"] + [str(v) for v in dims] +
                        ["-P"] + [str(v) for v in grid] + ["-solver", "1", "-rlx", str(rlx)],
                        capture_output=True, text=True, timeout=300, env=dict(os.environ, OMP_NUM_THREADS="1"))
    assert ij.returncode == 0, ij.stderr[-300:]
    its = int(re.search(r"^Iterations = (\d+)", ij.stdout, re.M).group(1))
    fin = float(re.search(r"Final Relative Residual Norm = (\S+)", ij.stdout).group(1))
    assert int(m.group(4)) == its
    assert abs(float(m.group(5)) - fin) <= 2e-6 * fin


def test_native_mode_has_no_cpu_fallback():
    """device path on, no GPU: every rank's solve raises the hypre error flag (0 iterations), nothing hangs"""
    try:
        gpu = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True, timeout=30).returncode == 0
    except Exception:
        gpu = False
    if gpu:
        pytest.skip("GPU present")
    argv = [sys.executable, os.path.join(ROOT, "tests", "native_worker.py"), "2", "1", "1", "8", "8", "8", "18"]
    out = launch(2, argv, extra_env={"NATIVE_WORKER_DEVICE": "1"}, timeout=120)
    assert all(rc == 0 for rc, _, _ in out.values()), {r: v[2][-400:] for r, v in out.items()}
    assert "iterations=0" in out[0][1]
    assert "no CPU fallback" in out[0][2] + out[1][2]
