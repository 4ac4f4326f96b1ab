"""CPU test of the text the binding prints for GMRES / BiCGSTAB solves (`ij -solver 3 / 4 / 9 / 10`).

hypre_glue.c interposes hypre_GMRESSolve / hypre_BiCGSTABSolve: the iterations run on the device, the residual
table and the "Final L2 norm of residual" line are printed by the binding from the history the device solve
returns (hypre_b200_krylov_report).  Here the reference solves on the CPU with the print level ij uses, its
output is captured, and the report function is fed the history the reference logged: both texts must agree
line by line (labels identical, numbers to the printed digits)."""
import ctypes as C

import numpy as np
import pytest

from test_amg_report_host import capture_stdout, same_text


@pytest.mark.parametrize("kind,precond", [("gmres", "amg"), ("gmres", "ds"), ("cogmres", "amg"), ("cogmres", "ds"), ("flexgmres", "amg"), ("flexgmres", "ds"), ("lgmres", "amg"), ("lgmres", "ds"),
                                          ("bicgstab", "amg"), ("bicgstab", "ds")])
def test_krylov_report_equals_the_reference_text(H, kind, precond):
    L = H.lib()
    L.hypre_b200_krylov_report.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_int]
    L.hypre_b200_krylov_report.restype = None
    H.set_device_path(False)
    try:
        A = H.laplacian(13, 12, 11)
        b, x = A.vector(1.0), A.vector(0.0)
        s = {"gmres": H.ParCSRGMRES, "cogmres": H.ParCSRCOGMRES, "flexgmres": H.ParCSRFlexGMRES,
             "lgmres": H.ParCSRLGMRES}[kind](print_level=3, max_iter=300) if kind != "bicgstab" \
            else H.ParCSRBiCGSTAB(print_level=3, logging=3, max_iter=300)
        if precond == "amg":
            amg = H.BoomerAMG(relax_type=18)
            s.SetPrecond(amg)
        else:
            s.SetDiagScalePrecond()
        s.Setup(A, b, x)
        H.clear_errors()
        ref_text = capture_stdout(lambda: s.Solve(A, b, x))
        assert H.error_flag() == 0 and s.converged == 1
    finally:
        H.set_device_path(True)
    norms = np.ascontiguousarray(s.norms())
    it = s.num_iterations
    b_norm = float(np.sqrt(b.numpy() @ b.numpy()))
    ours = capture_stdout(lambda: L.hypre_b200_krylov_report({"gmres": 0, "flexgmres": 0, "lgmres": 0, "bicgstab": 1, "cogmres": 2}[kind], norms.ctypes.data_as(C.c_void_p),
                                                             it, b_norm, s.final_rel_norm, 1))
    assert "Final L2 norm of residual" in ref_text and "Iters     resid.norm     conv.rate  rel.res.norm" in ref_text
    same_text(ours, ref_text)
    # blank lines included: the reference's drivers' goldens are compared as text
    assert ours.count("\n") == ref_text.count("\n")
