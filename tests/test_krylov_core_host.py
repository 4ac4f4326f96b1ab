"""CPU tests of the GMRES / BiCGSTAB loops the device library runs (hypre_b200/csrc/krylov_core.hpp).

The loops are templates over a vector backend; libb200amg.so instantiates them over its CUDA backend only
(krylov.cu).  Here the SAME header is instantiated over a host backend that lives under tests/
(tests/hostbackend/krylov_host_backend.cpp, test infrastructure: operator and preconditioner are the oracle's)
and every result is compared with orc_gmres / orc_bicgstab, which tests/test_oracle_krylov.py pins to the
compiled reference.  What this covers without a GPU: Givens rotations, restarts, the true-residual checks,
max_iter / breakdown / zero-residual / Inf-NaN exits, the logged norms.  What only the GPU tests cover: the
device functors (tests/test_gpu_zz_krylov.py)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest
import scipy.sparse as sp

import oracle_util as O
from hypre_b200 import ROOT

HB = os.path.join(ROOT, "tests", "hostbackend")


from hypre_b200.capi import BicgstabParams, GmresParams, KrylovResult, LgmresParams  # noqa: E402  (layout of include/b200amg.h)


@pytest.fixture(scope="module")
def kht(built):
    so = os.path.join(HB, "libkrylov_host_test.so")
    src = os.path.join(HB, "krylov_host_backend.cpp")
    hdr = os.path.join(ROOT, "hypre_b200", "csrc", "krylov_core.hpp")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        cmd = ["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-fPIC", "-shared", "-Wall", "-o", so, src,
               "-L" + os.path.join(ROOT, "oracle", "_ref"), "-loracle", "-Wl,-rpath," + os.path.join(ROOT, "oracle", "_ref")]
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    O.lib()
    L = C.CDLL(so)
    vp, i = C.c_void_p, C.c_int
    L.kht_gmres.argtypes = [i, vp, vp, vp, vp, i, vp, vp, vp, vp, vp]
    L.kht_bicgstab.argtypes = [i, vp, vp, vp, vp, i, vp, vp, vp, vp, vp]
    L.kht_lgmres.argtypes = [i, vp, vp, vp, vp, i, vp, vp, vp, vp, vp]
    return L


@pytest.fixture(scope="module")
def ref_off(H):
    H.set_device_path(False)
    yield H
    H.set_device_path(True)


def _one_level(Ad):
    return O.OracleAMG.from_arrays([dict(A=Ad, P=None)], [18, 18, 18, 18], [1, 1, 1, 1], 1, 0, 2, 1, None)


def run_core(kht, kind, Ad, orc, precond, b, x0, tol=1e-8, a_tol=0.0, k_dim=5, min_iter=0, max_iter=100, skip=0, cf_tol=0.0,
             stop_crit=0):
    ia, ja, aa = O.csr_arrays(Ad)
    x = np.array(x0, dtype=np.float64, copy=True)
    norms = np.zeros(max_iter + 1)
    res = KrylovResult()
    b = np.ascontiguousarray(b, dtype=np.float64)
    h = orc.h if orc is not None else None
    if kind in ("gmres", "cogmres", "flexgmres"):
        prm = GmresParams(tol, a_tol, max_iter, min_iter, k_dim, skip, 1 if precond == 1 else 0, 1 if kind == "cogmres" else 0,
                          1 if kind == "flexgmres" else 0, cf_tol)
        st = kht.kht_gmres(Ad.shape[0], O.p(ia), O.p(ja), O.p(aa), h, precond, C.byref(prm), O.p(b), O.p(x),
                           C.byref(res), O.p(norms))
    elif kind == "lgmres":
        prm = LgmresParams(tol, a_tol, max_iter, min_iter, k_dim, 2, 1, 1 if precond == 1 else 0, cf_tol)
        st = kht.kht_lgmres(Ad.shape[0], O.p(ia), O.p(ja), O.p(aa), h, precond, C.byref(prm), O.p(b), O.p(x),
                            C.byref(res), O.p(norms))
    else:
        prm = BicgstabParams(tol, a_tol, max_iter, min_iter, 1 if precond == 1 else 0, stop_crit, cf_tol)
        st = kht.kht_bicgstab(Ad.shape[0], O.p(ia), O.p(ja), O.p(aa), h, precond, C.byref(prm), O.p(b), O.p(x),
                              C.byref(res), O.p(norms))
    assert st == 0
    return x, res, norms


def check_against_oracle(kht, kind, Ad, orc, precond, b, x0, **kw):
    x, res, norms = run_core(kht, kind, Ad, orc, precond, b, x0, **kw)
    okw = dict(precond=precond, tol=kw.get("tol", 1e-8), a_tol=kw.get("a_tol", 0.0), min_iter=kw.get("min_iter", 0),
               max_iter=kw.get("max_iter", 100))
    if kind in ("gmres", "cogmres", "flexgmres", "lgmres"):
        okw.update(k_dim=kw.get("k_dim", 5), skip_real_r_check=kw.get("skip", 0))
    with O.krylov_options(cf_tol=kw.get("cf_tol", 0.0), stop_crit=kw.get("stop_crit", 0)):
        xo, it, no, rel, conv = getattr(orc, kind)(b, x0, **okw)
    assert res.num_iterations == it and res.converged == conv
    np.testing.assert_allclose(norms[:it + 1], no, rtol=1e-12)
    np.testing.assert_allclose(res.rel_residual_norm, rel, rtol=1e-12)
    np.testing.assert_allclose(x, xo, rtol=1e-12, atol=1e-14)
    return res


@pytest.mark.parametrize("kind", ["gmres", "cogmres", "flexgmres", "lgmres", "bicgstab"])
@pytest.mark.parametrize("name", ["lap7_24_rlx16", "lap27_20_rlx11", "lap7_24_rlx18_W", "lap7_24_rlx7_w08"])
def test_core_with_amg_preconditioner(kht, ref_off, kind, name):
    H = ref_off
    c = O.CASES[name]
    A = getattr(H, c["gen"])(*c["dims"])
    b, x = A.vector(1.0), A.vector(0.0)
    amg = H.BoomerAMG(relax_type=c["rlx"], **c.get("opts", {}))
    amg.Setup(A, b, x)
    orc = O.OracleAMG(amg)
    Ad, _ = A.blocks()
    res = check_against_oracle(kht, kind, Ad, orc, 0, b.numpy().copy(), np.zeros(x.size))
    assert res.converged == 1 and res.error_flag == 0 and 3 <= res.num_iterations <= 20
    orc.destroy()


@pytest.mark.parametrize("kind,k_dim,precond", [("gmres", 5, 1), ("gmres", 3, 1), ("gmres", 20, 1), ("gmres", 1, 1),
                                                ("gmres", 7, 2), ("cogmres", 5, 1), ("cogmres", 6, 2), ("cogmres", 13, 1),
                                                ("flexgmres", 5, 1), ("flexgmres", 4, 2), ("flexgmres", 11, 1),
                                                ("lgmres", 5, 1), ("lgmres", 3, 2), ("lgmres", 12, 1),
                                                ("bicgstab", 0, 1), ("bicgstab", 0, 2)])
def test_core_restarts_and_long_runs(kht, ref_off, kind, k_dim, precond):
    """diagonal scaling / no preconditioner: tens to hundreds of iterations, many restart cycles"""
    A = ref_off.laplacian(12, 11, 10)
    Ad, _ = A.blocks()
    orc = _one_level(Ad)
    b = 1.0 + 0.01 * np.cos(np.arange(Ad.shape[0]))
    res = check_against_oracle(kht, kind, Ad, orc, precond, b, np.zeros(Ad.shape[0]), k_dim=k_dim, max_iter=2000)
    assert res.converged == 1 and res.num_iterations > 15
    orc.destroy()


@pytest.mark.parametrize("kind", ["gmres", "cogmres", "flexgmres", "lgmres", "bicgstab"])
@pytest.mark.parametrize("max_iter", [0, 1, 7, 10])
def test_core_max_iter_exit_sets_conv_error(kht, ref_off, kind, max_iter):
    A = ref_off.laplacian(10, 9, 8)
    Ad, _ = A.blocks()
    orc = _one_level(Ad)
    x0 = 0.1 * np.sin(np.arange(Ad.shape[0]))
    res = check_against_oracle(kht, kind, Ad, orc, 1, np.ones(Ad.shape[0]), x0, k_dim=4, max_iter=max_iter)
    assert res.num_iterations == max_iter and res.converged == 0
    assert res.error_flag == 256      # iter >= max_iter && r_norm > epsilon (gmres.c:908, bicgstab.c:566), also for max_iter 0
    orc.destroy()


@pytest.mark.parametrize("kind", ["gmres", "cogmres", "flexgmres", "lgmres", "bicgstab"])
def test_core_min_iter_skip_check_and_abs_tol(kht, ref_off, kind):
    A = ref_off.laplacian(9, 8, 7)
    Ad, _ = A.blocks()
    orc = _one_level(Ad)
    n = Ad.shape[0]
    b = np.ones(n)
    check_against_oracle(kht, kind, Ad, orc, 1, b, np.zeros(n), tol=1e-2, min_iter=9, max_iter=200)
    check_against_oracle(kht, kind, Ad, orc, 1, b, np.zeros(n), tol=0.0, a_tol=1e-5, max_iter=200)
    if kind in ("gmres", "cogmres"):
        check_against_oracle(kht, kind, Ad, orc, 1, b, np.zeros(n), k_dim=6, skip=1, max_iter=200)
    orc.destroy()


@pytest.mark.parametrize("kind", ["gmres", "cogmres", "flexgmres", "lgmres", "bicgstab"])
def test_core_trivial_inputs(kht, ref_off, kind):
    """zero right-hand side with zero guess (r_norm == 0: immediate return), exact initial guess (already
    converged before the first iteration), Inf in b (error flag 1, as hypre_error(HYPRE_ERROR_GENERIC))"""
    A = ref_off.laplacian(6, 5, 4)
    Ad, _ = A.blocks()
    orc = _one_level(Ad)
    n = Ad.shape[0]
    x, res, norms = run_core(kht, kind, Ad, orc, 1, np.zeros(n), np.zeros(n))
    assert res.num_iterations == 0 and res.error_flag == 0 and not x.any()
    xs = np.cos(np.arange(n))
    b = Ad @ xs
    x, res, norms = run_core(kht, kind, Ad, orc, 1, b, xs)
    assert res.num_iterations == 0 and res.error_flag == 0
    np.testing.assert_array_equal(x, xs)
    xo, it, no, rel, conv = getattr(orc, kind)(b, xs, precond=1)
    assert it == 0 and conv == res.converged
    binf = np.ones(n)
    binf[3] = np.inf
    x, res, norms = run_core(kht, kind, Ad, orc, 1, binf, np.zeros(n))
    assert res.error_flag == 1 and res.num_iterations == 0
    orc.destroy()


def test_core_bicgstab_breakdown(kht, built):
    """<r0, A C p> == 0 in the first iteration: the reference raises "BiCGSTAB broke down!! divide by near zero"
    (bicgstab.c:433-437) and returns with num_iterations == 0 and x untouched"""
    Ad = sp.csr_matrix(np.array([[0.0, 1.0], [-1.0, 0.0]]))
    orc = _one_level(Ad)
    b = np.array([1.0, 0.0])
    x, res, norms = run_core(kht, "bicgstab", Ad, orc, 2, b, np.zeros(2))
    assert res.error_flag == 2 and res.num_iterations == 0 and res.converged == 0
    np.testing.assert_array_equal(x, np.zeros(2))
    xo, it, no, rel, conv = orc.bicgstab(b, np.zeros(2), precond=2)
    assert it == -2
    orc.destroy()


def test_product_library_does_not_contain_the_host_backend(built):
    """the host backend is test infrastructure: neither product library exports or links it"""
    from hypre_b200 import LIB_B200, LIB_HYPRE_B200
    for lib in (LIB_B200, LIB_HYPRE_B200):
        out = subprocess.run(["nm", "-D", lib], capture_output=True, text=True).stdout
        assert "kht_" not in out and "HostKrylovOps" not in out
        ldd = subprocess.run(["ldd", lib], capture_output=True, text=True).stdout
        assert "krylov_host_test" not in ldd
    # and the device library does not link the oracle
    assert "liboracle" not in subprocess.run(["ldd", LIB_B200], capture_output=True, text=True).stdout


@pytest.mark.parametrize("nranks", [2, 3])
def test_krylov_loops_on_row_partitioned_ranks_over_gloo(kht, nranks):
    """world_size 2 and 3 over gloo (tests/dist_krylov_worker.py): the same loops on a row-partitioned problem --
    halo exchange by point-to-point messages following b200_rowpart's plan, inner products by all_reduce -- reach
    the single-rank oracle's iteration counts, histories and solution for GMRES, COGMRES, FlexGMRES and BiCGSTAB"""
    import sys
    env = dict(os.environ, OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nranks),
           "--master-addr", "127.0.0.1", "--master-port", str(29550 + nranks), os.path.join(ROOT, "tests", "dist_krylov_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0 and "DIST_KRYLOV_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]


@pytest.mark.parametrize("kind", ["gmres", "cogmres", "flexgmres", "lgmres", "bicgstab"])
def test_core_convergence_factor_tolerance_and_stop_crit(kht, ref_off, kind):
    """cf_tol abandons a slowly converging solve (GMRES-type loops leave without the update of the running cycle);
    BiCGSTAB's stop_crit makes tol an absolute bound"""
    A = ref_off.laplacian(9, 8, 7)
    Ad, _ = A.blocks()
    orc = _one_level(Ad)
    n = Ad.shape[0]
    for cf in (0.7, 0.95):
        res = check_against_oracle(kht, kind, Ad, orc, 1, np.ones(n), np.zeros(n), k_dim=6, max_iter=300, cf_tol=cf)
        if cf == 0.7:
            assert res.converged == 0 and res.num_iterations < 20 and res.error_flag == 0
    if kind == "bicgstab":
        res = check_against_oracle(kht, kind, Ad, orc, 1, np.ones(n), np.zeros(n), tol=1e-3, max_iter=300, stop_crit=1)
        assert res.converged == 1
    orc.destroy()


