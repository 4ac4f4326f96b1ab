"""CPU tests: pin the plain-C restatements of hypre_GMRESSolve (krylov/gmres.c:274-930) and
hypre_BiCGSTABSolve (krylov/bicgstab.c:226-640) in oracle/amg_oracle_cycle.c against
 (a) the compiled reference library called in-process on the same hierarchy (full precision), and
 (b) the reference's own driver (oracle/_ref/ij_ref -solver 3/4/9/10): iteration counts and the printed
     final relative residual.
Tolerances are the north star's: residual norms 1e-9 relative, identical iteration counts."""
import os
import re
import subprocess

import numpy as np
import pytest

import oracle_util as O


@pytest.fixture(scope="module")
def ref_off(H):
    """reference CPU path, ONE OpenMP thread: the reductions of the reference then run in index order like the
    oracle's, and the two agree to 1e-9 over whole GMRES / BiCGSTAB histories.  (With several threads the
    partial sums differ in the last bit; the Arnoldi and BiCGSTAB recurrences carry that to ~1e-16 ||r_0||
    absolute, i.e. 3e-8 relative on a residual reduced by 1e-8 -- measured)"""
    import ctypes as C
    gomp = C.CDLL("libgomp.so.1")
    gomp.omp_get_max_threads.restype = C.c_int
    before = gomp.omp_get_max_threads()
    gomp.omp_set_num_threads(1)
    H.set_device_path(False)
    yield H
    H.set_device_path(True)
    gomp.omp_set_num_threads(before)


def _solver(H, kind, amg, k_dim=5, tol=1e-8, max_iter=100):
    if kind in ("gmres", "cogmres", "flexgmres", "lgmres"):
        cls = {"gmres": H.ParCSRGMRES, "cogmres": H.ParCSRCOGMRES, "flexgmres": H.ParCSRFlexGMRES,
               "lgmres": H.ParCSRLGMRES}[kind]
        s = cls(k_dim=k_dim, tol=tol, max_iter=max_iter, print_level=1)   # print_level 1: norms[] logged
    else:
        s = H.ParCSRBiCGSTAB(tol=tol, max_iter=max_iter, print_level=0, logging=1)
    if amg is not None:
        s.SetPrecond(amg)
    else:
        s.SetDiagScalePrecond()
    return s


def _converged(s):
    """hypre_*Data.converged; HYPRE_FlexGMRESGetConverged exists like the GMRES one"""
    return s.converged


def _case(H, name):
    c = O.CASES[name]
    A = getattr(H, c["gen"])(*c["dims"])
    b, x = A.vector(1.0), A.vector(0.0)
    amg = H.BoomerAMG(relax_type=c["rlx"], **c.get("opts", {}))
    return A, b, x, amg


@pytest.mark.parametrize("kind", ["gmres", "cogmres", "flexgmres", "lgmres", "bicgstab"])
@pytest.mark.parametrize("name", ["lap7_24_rlx16", "lap7_32_rlx18", "lap27_20_rlx11", "lap7_24_rlx18_W", "lap7_24_rlx18_pc3"])
def test_oracle_amg_krylov_matches_reference(ref_off, kind, name):
    H = ref_off
    A, b, x, amg = _case(H, name)
    s = _solver(H, kind, amg)
    s.Setup(A, b, x)
    H.clear_errors()
    s.Solve(A, b, x)
    assert H.error_flag() == 0
    orc = O.OracleAMG(amg)
    xo, it, norms, rel, conv = getattr(orc, kind)(b.numpy(), np.zeros(x.size))
    assert it == s.num_iterations and conv == _converged(s) == 1
    np.testing.assert_allclose(norms, s.norms(), rtol=1e-9)
    np.testing.assert_allclose(rel, s.final_rel_norm, rtol=1e-9)
    np.testing.assert_allclose(xo, x.numpy(), rtol=1e-9, atol=1e-12)
    orc.destroy()


@pytest.mark.parametrize("kind,k_dim", [("gmres", 5), ("gmres", 3), ("gmres", 20), ("cogmres", 5), ("cogmres", 9), ("cogmres", 20),
                                        ("flexgmres", 5), ("flexgmres", 8), ("lgmres", 5), ("lgmres", 9), ("lgmres", 3),
                                        ("bicgstab", 0)])
def test_oracle_diagscale_krylov_matches_reference(ref_off, kind, k_dim):
    """ij -solver 4 / 10: diagonal scaling, many iterations, several GMRES restarts"""
    H = ref_off
    A = H.laplacian(12, 11, 10)
    b, x = A.vector(1.0), A.vector(0.0)
    s = _solver(H, kind, None, k_dim=k_dim, max_iter=500)
    s.Setup(A, b, x)
    H.clear_errors()
    s.Solve(A, b, x)
    assert H.error_flag() == 0
    # the oracle only needs level 0 here: a one-level "hierarchy" built from A
    Ad, _ = A.blocks()
    orc = O.OracleAMG.from_arrays([dict(A=Ad, P=None)], [18, 18, 18, 18], [1, 1, 1, 1], 1, 0, 2, 1, None)
    kw = dict(precond=1, max_iter=500)
    if kind != "bicgstab":
        kw["k_dim"] = k_dim
    xo, it, norms, rel, conv = getattr(orc, kind)(b.numpy(), np.zeros(x.size), **kw)
    assert it == s.num_iterations and it > 10 and conv == _converged(s) == 1
    # long recurrences: rounding differences of the compiled reference (same operations, same order) stay tiny
    np.testing.assert_allclose(norms, s.norms(), rtol=1e-9)
    np.testing.assert_allclose(xo, x.numpy(), rtol=1e-9, atol=1e-12)
    orc.destroy()


def test_oracle_gmres_max_iter_and_nonzero_guess(ref_off):
    """max_iter reached in the middle of a restart cycle (HYPRE_ERROR_CONV), nonzero initial guess"""
    H = ref_off
    A = H.laplacian(10, 9, 8)
    b, x = A.vector(1.0), A.vector(0.0)
    x0 = np.sin(np.arange(x.size)) * 0.1
    x.numpy()[:] = x0
    s = _solver(H, "gmres", None, k_dim=4, max_iter=7)
    s.Setup(A, b, x)
    H.clear_errors()
    s.Solve(A, b, x)
    assert H.error_flag() & 256
    H.clear_errors()
    Ad, _ = A.blocks()
    orc = O.OracleAMG.from_arrays([dict(A=Ad, P=None)], [18, 18, 18, 18], [1, 1, 1, 1], 1, 0, 2, 1, None)
    xo, it, norms, rel, conv = orc.gmres(b.numpy(), x0, precond=1, k_dim=4, max_iter=7)
    assert it == s.num_iterations == 7 and conv == s.converged == 0
    np.testing.assert_allclose(norms, s.norms(), rtol=1e-9)
    np.testing.assert_allclose(rel, s.final_rel_norm, rtol=1e-9)
    np.testing.assert_allclose(xo, x.numpy(), rtol=1e-9, atol=1e-12)
    orc.destroy()


IJ_KRYLOV = {
    "gmres_amg": ("-laplacian -n 24 24 24 -solver 3 -rlx 16", "gmres", 0),
    "gmres_ds": ("-laplacian -n 12 11 10 -solver 4", "gmres", 1),
    "cogmres_amg": ("-laplacian -n 24 24 24 -solver 16 -rlx 18", "cogmres", 0),
    "cogmres_ds": ("-laplacian -n 12 11 10 -solver 17", "cogmres", 1),
    "flexgmres_amg": ("-laplacian -n 24 24 24 -solver 61 -rlx 18", "flexgmres", 0),
    "flexgmres_ds": ("-laplacian -n 12 11 10 -solver 60", "flexgmres", 1),
    "lgmres_amg": ("-laplacian -n 24 24 24 -solver 51 -rlx 18", "lgmres", 0),
    "lgmres_ds": ("-laplacian -n 12 11 10 -solver 50", "lgmres", 1),
    "bicgstab_amg": ("-27pt -n 20 20 20 -solver 9 -rlx 11", "bicgstab", 0),
    "bicgstab_ds": ("-laplacian -n 12 11 10 -solver 10", "bicgstab", 1),
}


@pytest.mark.parametrize("key", sorted(IJ_KRYLOV))
def test_oracle_krylov_matches_ij_driver(ref_off, key):
    """the reference's own driver on the same problem: iteration count and printed final residual"""
    from hypre_b200 import ROOT
    H = ref_off
    ij = os.path.join(ROOT, "oracle", "_ref", "ij_ref")
    if not os.path.exists(ij):
        pytest.skip("ij_ref not built")
    args, kind, precond = IJ_KRYLOV[key]
    r = subprocess.run([ij] + args.split(), capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, OMP_NUM_THREADS="1"))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    it_ref = int(re.search(r"Iterations = (\d+)", r.stdout).group(1))
    fin_ref = float(re.search(r"Final.*Relative Residual Norm = (\S+)", r.stdout).group(1))
    a = args.split()
    dims = tuple(int(v) for v in a[a.index("-n") + 1:a.index("-n") + 4])
    A = H.laplacian_27pt(*dims) if "-27pt" in a else H.laplacian(*dims)
    b, x = A.vector(1.0), A.vector(0.0)
    if precond == 0:
        amg = H.BoomerAMG(relax_type=int(a[a.index("-rlx") + 1]))
        amg.Setup(A, b, x)
        orc = O.OracleAMG(amg)
    else:
        Ad, _ = A.blocks()
        orc = O.OracleAMG.from_arrays([dict(A=Ad, P=None)], [18, 18, 18, 18], [1, 1, 1, 1], 1, 0, 2, 1, None)
    xo, it, norms, rel, conv = getattr(orc, kind)(b.numpy(), np.zeros(x.size), precond=precond, max_iter=1000)
    assert it == it_ref
    np.testing.assert_allclose(rel, fin_ref, rtol=2e-6)
    orc.destroy()


@pytest.mark.parametrize("kind", ["gmres", "bicgstab"])
@pytest.mark.parametrize("opts", [dict(tol=1e-2, min_iter=9), dict(tol=0.0, a_tol=1e-5), dict(max_iter=0),
                                  dict(skip_real_r_check=1, k_dim=6)])
def test_oracle_krylov_options_match_reference(ref_off, kind, opts):
    """min_iter, absolute tolerance, max_iter 0, skip_real_r_check (GMRES only)"""
    H = ref_off
    if kind == "bicgstab" and "skip_real_r_check" in opts:
        pytest.skip("GMRES option")
    A = H.laplacian(9, 8, 7)
    b, x = A.vector(1.0), A.vector(0.0)
    s = _solver(H, kind, None, k_dim=opts.get("k_dim", 5), tol=opts.get("tol", 1e-8), max_iter=opts.get("max_iter", 200))
    if "min_iter" in opts:
        s.SetMinIter(opts["min_iter"])
    if "a_tol" in opts:
        s.SetAbsoluteTol(float(opts["a_tol"]))
    if "skip_real_r_check" in opts:
        s.SetSkipRealResidualCheck(1)
    s.Setup(A, b, x)
    H.clear_errors()
    s.Solve(A, b, x)
    flag = H.error_flag()
    H.clear_errors()
    Ad, _ = A.blocks()
    orc = O.OracleAMG.from_arrays([dict(A=Ad, P=None)], [18, 18, 18, 18], [1, 1, 1, 1], 1, 0, 2, 1, None)
    kw = dict(precond=1, tol=opts.get("tol", 1e-8), a_tol=opts.get("a_tol", 0.0), min_iter=opts.get("min_iter", 0),
              max_iter=opts.get("max_iter", 200))
    if kind == "gmres":
        kw.update(k_dim=opts.get("k_dim", 5), skip_real_r_check=opts.get("skip_real_r_check", 0))
    xo, it, norms, rel, conv = getattr(orc, kind)(b.numpy(), np.zeros(x.size), **kw)
    assert it == s.num_iterations and conv == s.converged
    assert flag == (256 if opts.get("max_iter", 200) == 0 else 0)
    if it > 0:
        np.testing.assert_allclose(norms, s.norms(), rtol=1e-9)
    np.testing.assert_allclose(rel, s.final_rel_norm, rtol=1e-9)
    np.testing.assert_allclose(xo, x.numpy(), rtol=1e-9, atol=1e-12)
    orc.destroy()


@pytest.mark.parametrize("kind", ["gmres", "cogmres", "flexgmres", "lgmres", "bicgstab"])
@pytest.mark.parametrize("cf_tol", [0.7, 0.95])
def test_oracle_convergence_factor_tolerance_matches_reference(ref_off, kind, cf_tol):
    """cf_tol (HYPRE_*SetConvergenceFactorTol): the solve is abandoned when the weighted average convergence factor
    exceeds it -- GMRES-type solvers then leave WITHOUT the update of the running restart cycle (gmres.c:732-735)"""
    H = ref_off
    A = H.laplacian(9, 8, 7)
    b, x = A.vector(1.0), A.vector(0.0)
    s = _solver(H, kind, None, k_dim=6, max_iter=300)
    s.SetConvergenceFactorTol(float(cf_tol))
    s.Setup(A, b, x)
    H.clear_errors()
    s.Solve(A, b, x)
    H.clear_errors()
    Ad, _ = A.blocks()
    orc = O.OracleAMG.from_arrays([dict(A=Ad, P=None)], [18, 18, 18, 18], [1, 1, 1, 1], 1, 0, 2, 1, None)
    kw = dict(precond=1, max_iter=300)
    if kind != "bicgstab":
        kw["k_dim"] = 6
    with O.krylov_options(cf_tol=cf_tol):
        xo, it, norms, rel, conv = getattr(orc, kind)(b.numpy(), np.zeros(x.size), **kw)
    assert it == s.num_iterations and conv == s.converged
    if cf_tol == 0.7:
        assert it < 20 and conv == 0          # diagonal scaling on a Laplacian contracts by ~0.9 per step: abandoned
    np.testing.assert_allclose(norms, s.norms(), rtol=1e-9)
    np.testing.assert_allclose(rel, s.final_rel_norm, rtol=1e-9)
    np.testing.assert_allclose(xo, x.numpy(), rtol=1e-9, atol=1e-12)
    orc.destroy()


def test_oracle_bicgstab_stop_crit_matches_reference(ref_off):
    """stop_crit = 1: |r| <= tol as an ABSOLUTE bound (a_tol when that is set), bicgstab.c:372-383"""
    H = ref_off
    A = H.laplacian(9, 8, 7)
    Ad, _ = A.blocks()
    orc = O.OracleAMG.from_arrays([dict(A=Ad, P=None)], [18, 18, 18, 18], [1, 1, 1, 1], 1, 0, 2, 1, None)
    for tol, a_tol in ((1e-3, 0.0), (1e-3, 1e-6)):
        b, x = A.vector(1.0), A.vector(0.0)
        s = _solver(H, "bicgstab", None, tol=tol, max_iter=300)
        s.SetStopCrit(1)
        s.SetAbsoluteTol(float(a_tol))
        s.Setup(A, b, x)
        H.clear_errors()
        s.Solve(A, b, x)
        assert H.error_flag() == 0
        with O.krylov_options(stop_crit=1):
            xo, it, norms, rel, conv = orc.bicgstab(b.numpy(), np.zeros(x.size), precond=1, tol=tol, a_tol=a_tol, max_iter=300)
        assert it == s.num_iterations and conv == s.converged == 1
        np.testing.assert_allclose(norms, s.norms(), rtol=1e-9)
        np.testing.assert_allclose(xo, x.numpy(), rtol=1e-9, atol=1e-12)
    orc.destroy()


