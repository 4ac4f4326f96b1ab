"""CPU test of the text the binding prints for a stand-alone BoomerAMG solve (`ij -solver 0`).

hypre_glue.c interposes hypre_BoomerAMGSolve (src/parcsr_ls/par_amg_solve.c:21-424): the cycles run on the
device, the residual table and the closing statistics (average convergence factor, grid / operator / cycle
complexity -- the numbers TEST_ij/*.saved hold for -solver 0 jobs) are printed by the binding.  Here the
reference solves on the CPU with print level 3, its table is parsed, and the binding's report function is fed
the same history: both texts must agree line by line (labels identical, numbers to the printed digits).
The cycle complexity also checks b200_cycle_sweep_counts (V, W and F cycles, 1 and 2 sweeps) against
hypre_BoomerAMGCycle's own count."""
import ctypes as C
import os
import re
import tempfile

import numpy as np
import pytest

from hypre_b200 import capi

ROW = re.compile(r"^\s+Cycle\s+\d+\s+\S+e[-+]\d+")
NUM = re.compile(r"^[-+]?(\d+\.\d*|\.\d+|\d+)([eE][-+]?\d+)?$")


def capture_stdout(fn):
    """C-level stdout of the calling process while fn runs"""
    libc = C.CDLL(None)
    libc.fflush(None)
    saved = os.dup(1)
    with tempfile.TemporaryFile(mode="w+b") as tmp:
        os.dup2(tmp.fileno(), 1)
        try:
            fn()
            libc.fflush(None)
        finally:
            os.dup2(saved, 1)
            os.close(saved)
        tmp.seek(0)
        return tmp.read().decode()


def same_text(a, b, rtol=5e-6):
    la = [l.split() for l in a.splitlines() if l.strip()]
    lb = [l.split() for l in b.splitlines() if l.strip()]
    assert len(la) == len(lb), (a, b)
    for ta, tb in zip(la, lb):
        assert len(ta) == len(tb), (ta, tb)
        for x, y in zip(ta, tb):
            if NUM.match(x) and NUM.match(y):
                assert abs(float(x) - float(y)) <= rtol * abs(float(y)) + 1e-6 * (("e" not in y.lower())), (ta, tb)
            else:
                assert x == y, (ta, tb)


@pytest.mark.parametrize("cycle_type,fcycle,sweeps,relax", [(1, 0, 1, 18), (2, 0, 1, 18), (1, 1, 1, 16), (1, 0, 2, 18),
                                                            (2, 0, 2, 7)])
def test_report_equals_the_reference_text(H, cycle_type, fcycle, sweeps, relax):
    L = H.lib()
    L.hypre_b200_amg_report.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_double, C.c_int]
    L.hypre_b200_amg_report.restype = None
    A = H.laplacian(13, 12, 11)
    b = A.vector(1.0); x = A.vector(0.0)
    amg = H.BoomerAMG(relax_type=relax, as_precond=False)
    amg.SetCycleType(cycle_type); amg.SetFCycle(fcycle); amg.SetNumSweeps(sweeps)
    amg.SetTol(1e-8); amg.SetMaxIter(40); amg.SetPrintLevel(3)
    H.set_device_path(False)
    try:
        capture_stdout(lambda: amg.Setup(A, b, x))
        out = capture_stdout(lambda: amg.Solve(A, b, x))      # through the interposer, forwarded to the reference
    finally:
        H.set_device_path(True)
        H.lib().HYPRE_ClearAllErrors()
    lines = out.splitlines()
    init = [l for l in lines if l.strip().startswith("Initial")]
    rows = [l for l in lines if ROW.match(l)]
    assert len(init) == 1 and len(rows) >= 3
    hist = np.array([float(init[0].split()[1])] + [float(r.split()[2]) for r in rows])
    rhs_norm = float(np.sqrt(A.global_rows))
    assert abs(float(init[0].split()[2]) - hist[0] / rhs_norm) < 1e-6
    tail = out[out.index(rows[0]):]
    mine = capture_stdout(lambda: L.hypre_b200_amg_report(amg.h, len(rows), capi.ptr(hist), rhs_norm, int(len(rows) == 40)))
    same_text(mine, tail)
    assert "Average Convergence Factor" in mine and "cycle =" in mine
    amg.destroy(); A.destroy()


def test_not_converged_note(H):
    L = H.lib()
    L.hypre_b200_amg_report.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_double, C.c_int]
    L.hypre_b200_amg_report.restype = None
    A = H.laplacian(10, 10, 10)
    b = A.vector(1.0); x = A.vector(0.0)
    amg = H.BoomerAMG(relax_type=18, as_precond=False)
    amg.SetTol(1e-12); amg.SetMaxIter(2); amg.SetPrintLevel(3)
    H.set_device_path(False)
    try:
        capture_stdout(lambda: amg.Setup(A, b, x))
        out = capture_stdout(lambda: amg.Solve(A, b, x))
    finally:
        H.set_device_path(True)
        H.lib().HYPRE_ClearAllErrors()
    rows = [l for l in out.splitlines() if ROW.match(l)]
    init = [l for l in out.splitlines() if l.strip().startswith("Initial")][0]
    hist = np.array([float(init.split()[1])] + [float(r.split()[2]) for r in rows])
    mine = capture_stdout(lambda: L.hypre_b200_amg_report(amg.h, 2, capi.ptr(hist), float(np.sqrt(1000.0)), 1))
    same_text(mine, out[out.index(rows[0]):])
    assert "Convergence tolerance was not achieved" in mine
    amg.destroy(); A.destroy()


@pytest.mark.parametrize("nl,cycle_type,fcycle,sweeps,want", [
    (1, 1, 0, [3, 1, 1, 1], [3]),
    (4, 1, 0, [1, 1, 1, 1], [2, 2, 2, 1]),                  # V(1,1)
    (4, 1, 0, [1, 2, 1, 3], [3, 3, 3, 3]),                  # V(2,1), 3 sweeps on the coarsest
    (3, 2, 0, [1, 1, 1, 1], [2, 3, 2]),                     # W: the post-sweep of a visit is the pre-sweep of the next
    (4, 2, 0, [1, 1, 1, 1], [2, 3, 6, 4]),
    (4, 1, 1, [1, 1, 1, 1], [3, 5, 7, 4]),                  # F
])
def test_sweep_counts(nl, cycle_type, fcycle, sweeps, want):
    out = (C.c_longlong * nl)()
    s = (C.c_int * 4)(*sweeps)
    capi.check(capi.lib().b200_cycle_sweep_counts(nl, cycle_type, fcycle, s, out))
    assert list(out) == want
