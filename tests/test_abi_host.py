"""CPU tests of the boundary: the C-ABI library loads and exports every symbol that
include/b200amg.h declares, fails loudly without a GPU (no CPU fallback), and the host-side
mirror of the HYPRE API drives the reference's setup exactly like src/test/ij.c."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import oracle_util as O
from hypre_b200 import IJ_B200, LIB_B200, capi


def has_gpu():
    ctx = C.c_void_p()
    ok = capi.lib().b200_ctx_create(0, C.byref(ctx)) == 0
    if ok:
        capi.lib().b200_ctx_destroy(ctx)
    return ok


def test_every_declared_symbol_is_exported(built):
    names = capi.declared_symbols()
    assert len(names) >= 45
    L = C.CDLL(LIB_B200)
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    # and the ctypes table covers the whole header
    capi.lib()
    out = subprocess.run(["nm", "-D", "--defined-only", LIB_B200], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l and "b200_" in l}
    assert set(names) <= exported


def test_cuda_library_is_sm100a_only(built):
    out = subprocess.run(["cuobjdump", "-lelf", LIB_B200], capture_output=True, text=True).stdout
    archs = {l.split(".")[-2] for l in out.splitlines() if ".cubin" in l}
    assert archs == {"sm_100a"}, archs


def test_no_cpu_fallback_without_gpu(H):
    if has_gpu():
        pytest.skip("GPU present")
    L = capi.lib()
    ctx = C.c_void_p()
    assert L.b200_ctx_create(0, C.byref(ctx)) == -1          # B200_ERR_CUDA
    assert b"no CPU fallback" in L.b200_last_error()
    # through the HYPRE API: the solve raises the hypre error flag instead of running on the CPU
    H.set_device_path(True)
    A, b, x, amg, pcg = O.build_case(H, "lap7_24_rlx16")
    H.clear_errors()
    pcg.Solve(A, b, x)
    assert H.error_flag() & 1                                  # HYPRE_ERROR_GENERIC
    assert pcg.num_iterations == 0
    assert not np.any(x.numpy())                               # untouched
    H.clear_errors()


def test_standalone_amg_has_no_cpu_fallback_either(H):
    """HYPRE_BoomerAMGSolve (ij -solver 0) is interposed too: without a GPU it raises the error flag and
    leaves x alone; with the device path switched off (test hook) the same call reaches the reference"""
    if has_gpu():
        pytest.skip("GPU present")
    A = H.laplacian(8, 8, 8)
    b = A.vector(1.0); x = A.vector(0.0)
    amg = H.BoomerAMG(relax_type=18, as_precond=False)
    amg.SetTol(1e-8); amg.SetMaxIter(20)
    H.set_device_path(False)
    try:
        amg.Setup(A, b, x)
    finally:
        H.set_device_path(True)
    H.clear_errors()
    amg.Solve(A, b, x)
    assert H.error_flag() & 1 and not np.any(x.numpy())
    H.clear_errors()
    H.set_device_path(False)
    try:
        amg.Solve(A, b, x)
    finally:
        H.set_device_path(True)
    assert H.error_flag() == 0 and np.any(x.numpy())
    it = C.c_int()
    H.lib().HYPRE_BoomerAMGGetNumIterations(amg.h, C.byref(it))
    assert 3 < it.value < 20
    amg.destroy(); A.destroy()


def test_ij_driver_links_against_the_binding(built):
    """the reference's unchanged ij.c resolves hypre_PCGSolve to the interposer"""
    if not os.path.exists(IJ_B200):
        pytest.skip("ij_b200 not built")
    env = dict(os.environ, LD_DEBUG="bindings")
    r = subprocess.run([IJ_B200, "-laplacian", "-n", "6", "6", "6", "-solver", "1", "-rlx", "18"],
                       capture_output=True, text=True, env=env)
    lines = [l for l in r.stderr.splitlines() if "`hypre_PCGSolve'" in l and "binding file" in l]
    # calls made from inside the reference library (HYPRE_PCGSolve -> hypre_PCGSolve) land in the binding;
    # the binding's own RTLD_NEXT lookup is the only one that resolves to the reference
    from_ref = [l for l in lines if "libHYPRE_host.so" in l.split(" to ")[0]]
    assert from_ref and all("libHYPRE_b200.so" in l.split(" to ")[1] for l in from_ref), lines[:3]


def test_python_mirror_matches_ij_defaults(H):
    H.set_device_path(False)
    try:
        A, b, x, amg, pcg = O.build_case(H, "lap7_24_rlx16")
        prm = amg.params()
        assert prm["relax_type"][1:3] == [16, 16] and prm["relax_type"][3] in (9, 16)
        assert prm["cheby_order"] == 2 and prm["cheby_scale"] == 1 and prm["relax_order"] == 0
        g = O.golden()["lap7_24_rlx16"]
        assert amg.num_levels == g["num_levels"]
        pcg.Solve(A, b, x)
        assert pcg.num_iterations == g["iterations"]
        np.testing.assert_allclose(pcg.rel_norms(), g["rel_norms"], rtol=2e-6)
        # hierarchy accessors
        Ad, Ao = amg.A(1).blocks()
        Pd, _ = amg.P(0).blocks()
        assert Ao.shape[1] == 0 and Pd.shape == (24 ** 3, Ad.shape[0])
        dense = amg.coarse_dense()
        assert dense is None or dense.shape[0] == amg.A(amg.num_levels - 1).blocks()[0].shape[0]
    finally:
        H.set_device_path(True)


def test_unsupported_pcg_options_are_rejected_loudly(H):
    H.set_device_path(True)
    A, b, x, amg, pcg = O.build_case(H, "lap7_24_rlx16")
    pcg.SetRelChange(1)
    H.clear_errors()
    pcg.Solve(A, b, x)
    assert H.error_flag() & 1
    H.clear_errors()


def test_multirank_binding_fails_loudly_without_gpu(built):
    """native multi-rank mode (ij_b200_mp over hypre_b200/host/mpishim): without a device every rank reports the
    failure and the solve does not silently run on the CPU (0 iterations, hypre error raised)"""
    if has_gpu():
        pytest.skip("GPU present")
    from hypre_b200 import ROOT
    shim = os.path.join(ROOT, "hypre_b200", "lib", "mpirun_shim")
    ij_mp = os.path.join(ROOT, "hypre_b200", "lib", "ij_b200_mp")
    if not (os.path.exists(shim) and os.path.exists(ij_mp)):
        pytest.skip("multi-rank builds absent")
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run([shim, "-np", "2", ij_mp, "-n", "12", "12", "12", "-P", "2", "1", "1", "-solver", "1", "-rlx", "18"],
                       capture_output=True, text=True, timeout=120, env=env)
    err = r.stderr
    assert err.count("cannot create the distributed context") == 2, err[-800:]
    assert "no CPU fallback" in err
    assert "hierarchy upload failed" in err
    assert "Iterations = 0" in r.stdout


@pytest.mark.parametrize("kind", ["gmres", "cogmres", "flexgmres", "lgmres", "bicgstab"])
def test_gmres_bicgstab_have_no_cpu_fallback_and_reject_unsupported_options(H, kind):
    """hypre_GMRESSolve / hypre_BiCGSTABSolve are interposed: without a GPU the solve raises the hypre error flag
    and leaves x alone; options outside the device path are rejected the same way (never run on the CPU)"""
    if has_gpu():
        pytest.skip("GPU present")
    H.set_device_path(True)
    A = H.laplacian(8, 8, 8)
    b = A.vector(1.0); x = A.vector(0.0)
    s = {"gmres": H.ParCSRGMRES, "cogmres": H.ParCSRCOGMRES, "flexgmres": H.ParCSRFlexGMRES,
         "lgmres": H.ParCSRLGMRES, "bicgstab": H.ParCSRBiCGSTAB}[kind]()
    s.SetDiagScalePrecond()
    s.Setup(A, b, x)
    H.clear_errors()
    s.Solve(A, b, x)
    assert H.error_flag() & 1 and s.num_iterations == 0 and not np.any(x.numpy())
    H.clear_errors()
    if kind == "gmres":
        s.SetRelChange(1)
    elif kind == "cogmres":
        s.SetCGS(2)
    elif kind == "lgmres":
        s.SetKDim(2)                      # aug_dim (2) >= k_dim
    elif kind == "flexgmres":
        cb = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_double)(lambda data, it, rel: 0)
        H.lib().HYPRE_FlexGMRESSetModifyPC(s.h, cb)      # a callback that could retune the preconditioner
    else:
        pass                              # BiCGSTAB: stop_crit and cf_tol are implemented; nothing left to reject here
    s.Solve(A, b, x)
    assert H.error_flag() & 1 and not np.any(x.numpy())
    H.clear_errors()
    # the transposed product, too
    y = A.vector(0.0, cols=True)
    H.lib().HYPRE_ParCSRMatrixMatvecT(1.0, A.h, b.h, 0.0, y.h)
    assert H.error_flag() & 1 and not np.any(y.numpy())
    H.clear_errors()


def test_device_vector_bindings_table_and_mixed_operands(H):
    """hypre_b200_vector_bind_device: bind / rebind / unbind; syncing an unbound vector fails; with a
    half-bound pair of operands no solve runs (here, without a device, the upload fails first -- the rejection of
    mixed operands itself is asserted on the GPU, tests/test_gpu_zz_krylov.py).  The fake pointers are never
    dereferenced."""
    H.set_device_path(True)
    A = H.laplacian(6, 6, 6)
    b = A.vector(1.0); x = A.vector(0.0)
    assert b.bound_device() is None
    fake1, fake2 = C.c_void_p(0x1000), C.c_void_p(0x2000)
    assert b.bind_device(fake1) == 0 and b.bound_device() == 0x1000
    assert b.bind_device(fake2) == 0 and b.bound_device() == 0x2000       # rebinding replaces
    assert x.bound_device() is None
    s = H.ParCSRPCG()
    s.SetDiagScalePrecond()
    s.Setup(A, b, x)
    for solver in (s, H.ParCSRGMRES(), H.ParCSRBiCGSTAB()):
        if solver is not s:
            solver.SetDiagScalePrecond()
            solver.Setup(A, b, x)
        H.clear_errors()
        solver.Solve(A, b, x)
        assert H.error_flag() & 1 and not np.any(x.numpy())
    H.clear_errors()
    assert b.bind_device(None) == 0 and b.bound_device() is None         # unbound again
    assert b.bind_device(None) == 0                                     # idempotent
    assert b.sync_to_host() == -1 and b.sync_to_device() == -1          # not bound


def test_cycle_exchange_counts():
    """b200_cycle_exchange_counts: halo exchanges per level of one cycle (host arithmetic), against hand counts
    for V-cycles and against b200_cycle_sweep_counts for W / F cycles"""
    L = capi.lib()

    def counts(nl, cycle_type, fcycle, rt, ns, order, zg):
        rt, ns = np.array(rt, dtype=np.int32), np.array(ns, dtype=np.int32)
        a, p, r = (np.zeros(nl, dtype=np.int64) for _ in range(3))
        capi.check(L.b200_cycle_exchange_counts(nl, cycle_type, fcycle, capi.ptr(rt), capi.ptr(ns), order, zg,
                                                capi.ptr(a), capi.ptr(p), capi.ptr(r)))
        sw = np.zeros(nl, dtype=np.int64)
        capi.check(L.b200_cycle_sweep_counts(nl, cycle_type, fcycle, capi.ptr(ns), capi.ptr(sw)))
        return a.tolist(), p.tolist(), r.tolist(), sw.tolist()
    # V-cycle, l1-Jacobi, zero initial guess: the first pre-sweep of every level is u = w f ./ l1 (no SpMV) and
    # stays out; residual, post-sweep -> 2 per level; nothing on the dense coarsest level
    a, p, r, sw = counts(3, 1, 0, [0, 18, 18, 9], [1, 1, 1, 1], 2, 1)
    assert a == [2, 2, 0] and p == [1, 1, 0] and r == [1, 1, 0]
    # nonzero guess on level 0 only: its pre-sweep needs the halo, too
    a, p, r, sw = counts(3, 1, 0, [0, 18, 18, 9], [1, 1, 1, 1], 2, 0)
    assert a == [3, 2, 0]
    # Chebyshev order 2 = 2 SpMVs per sweep, 1 from a zero guess: 1 + residual + 2
    a, p, r, sw = counts(4, 1, 0, [0, 16, 16, 9], [1, 1, 1, 1], 2, 1)
    assert a == [4, 4, 4, 0] and p == r == [1, 1, 1, 0]
    # two-stage GS, two sweeps: (0 + 1) pre, residual, 2 post
    a, p, r, sw = counts(3, 1, 0, [0, 11, 11, 9], [2, 2, 2, 1], 2, 1)
    assert a == [4, 4, 0]
    # W and F cycles: every descent comes back (fwdP == revP), level l is entered revP[l-1] times, and with
    # Jacobi one sweep each: exchanges = sweeps - zero-guess first sweeps + residuals
    for ct, fc in ((2, 0), (1, 1), (2, 1)):
        a, p, r, sw = counts(5, ct, fc, [0, 18, 18, 9], [1, 1, 1, 1], 2, 1)
        assert p == r and r[-1] == 0
        entered = [1] + r[:-1]
        for l in range(4):
            assert a[l] == sw[l] - entered[l] + r[l], (ct, fc, l, a, sw, r)
        assert a[4] == 0


def test_device_binding_goes_away_with_the_vector(H):
    """hypre_ParVectorDestroy is interposed: a destroyed vector loses its device binding (its address may be reused)"""
    A = H.laplacian(5, 5, 5)
    v = A.vector(1.0)
    handle = v.h.value
    v.bind_device(C.c_void_p(0x3000))
    assert v.bound_device() == 0x3000
    v.destroy()
    L = H.lib()
    L.hypre_b200_vector_device.argtypes = [C.c_void_p]
    L.hypre_b200_vector_device.restype = C.c_void_p
    assert L.hypre_b200_vector_device(C.c_void_p(handle)) is None




def test_product_libraries_do_not_depend_on_the_oracle(built):
    """nothing under hypre_b200/lib links, rpaths or loads anything under oracle/ (the checker is built
    separately under oracle/_ref; the product's host library is hypre_b200/lib/libHYPRE_host.so)"""
    from hypre_b200 import LIB_DIR
    for name in sorted(os.listdir(LIB_DIR)):
        path = os.path.join(LIB_DIR, name)
        if not (name.endswith(".so") or name.startswith("ij_b200") or name == "mpirun_shim"):
            continue
        r = subprocess.run(["ldd", path], capture_output=True, text=True)
        assert "oracle" not in r.stdout, (name, r.stdout)
        d = subprocess.run(["readelf", "-d", path], capture_output=True, text=True).stdout
        assert "oracle" not in d, (name, d)
