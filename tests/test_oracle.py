"""CPU tests: pin the plain-C oracle (oracle/amg_oracle*.c) against
 (a) the golden PCG histories produced by the reference's own ij driver (tests/golden), and
 (b) the compiled reference library called in-process on the same hierarchy.
Tolerances are the north star's: SpMV 1e-12 relative per row, residual norms 1e-9 relative,
iteration counts identical."""
import ctypes as C

import numpy as np
import pytest

import oracle_util as O


def rel_rows(y, yref):
    scale = np.maximum(np.abs(yref), 1e-300)
    return np.max(np.abs(y - yref) / np.maximum(scale, np.max(np.abs(yref)) * 1e-3))


@pytest.fixture(scope="module")
def ref_off(H):
    H.set_device_path(False)
    yield H
    H.set_device_path(True)


@pytest.mark.parametrize("name", ["lap7_32_rlx18", "lap7_24_rlx16", "lap27_20_rlx11", "lap27_20_rlx12"])
def test_oracle_pcg_matches_golden(ref_off, name):
    H = ref_off
    g = O.golden()[name]
    A, b, x, amg, pcg = O.build_case(H, name)
    assert amg.num_levels == g["num_levels"]
    orc = O.OracleAMG(amg)
    xs, it, rel, conv = orc.pcg(b.numpy(), x.numpy())
    assert it == g["iterations"] and conv == 1
    # goldens are printed with 7 significant digits
    np.testing.assert_allclose(rel, g["rel_norms"], rtol=2e-6)
    # and the reference library in-process, to full precision
    pcg.Solve(A, b, x)
    assert pcg.num_iterations == it
    np.testing.assert_allclose(rel, pcg.rel_norms(), rtol=1e-9)
    np.testing.assert_allclose(xs, x.numpy(), rtol=1e-9, atol=1e-12)
    orc.destroy()


def test_python_driver_reproduces_ij_golden_c1(ref_off):
    """BASELINE.json configs[0]: ij -laplacian -n 100 100 100 -solver 1 -rlx 18 (13 iterations)."""
    H = ref_off
    g = O.golden()["lap7_100_rlx18"]
    A, b, x, amg, pcg = O.build_case(H, "lap7_100_rlx18")
    pcg.Solve(A, b, x)
    assert pcg.num_iterations == g["iterations"] == 13
    np.testing.assert_allclose(pcg.rel_norms(), g["rel_norms"], rtol=2e-6)
    assert amg.num_levels == 9


def test_oracle_matvec_and_transpose_vs_reference(ref_off):
    H = ref_off
    L = H.lib()
    A = H.laplacian_27pt(12, 11, 10)
    Ad, _ = A.blocks()
    n = Ad.shape[0]
    rng = np.random.default_rng(1)
    x = A.vector(0.0); b = A.vector(0.0); y = A.vector(0.0)
    x.numpy()[:] = rng.standard_normal(n)
    b.numpy()[:] = rng.standard_normal(n)
    for alpha, beta in [(1.0, 0.0), (-1.0, 1.0), (2.5, -0.75), (1.0, 1.0)]:
        L.hypre_ParCSRMatrixMatvecOutOfPlace(alpha, A.h, x.h, beta, b.h, y.h)
        yo = O.matvec(Ad, alpha, x.numpy(), beta, b.numpy())
        assert rel_rows(yo, y.numpy()) < 1e-12
        y.numpy()[:] = b.numpy()
        L.hypre_ParCSRMatrixMatvecT(alpha, A.h, x.h, beta, y.h)
        yt = O.matvecT(Ad, alpha, x.numpy(), beta, b.numpy())
        assert rel_rows(yt, y.numpy()) < 1e-12


@pytest.mark.parametrize("name,rtype", [("lap7_32_rlx18", 18), ("lap27_20_rlx11", 11), ("lap27_20_rlx12", 12)])
def test_oracle_relax_vs_reference(ref_off, name, rtype):
    """hypre_BoomerAMGRelax on every level of a real hierarchy vs the restatement."""
    H = ref_off
    L = H.lib()
    A, b, x, amg, pcg = O.build_case(H, name)
    rng = np.random.default_rng(2)
    for lev in range(amg.num_levels - 1):
        Al = amg.A(lev)
        Ad, _ = Al.blocks()
        n = Ad.shape[0]
        f = Al.vector(0.0); u = Al.vector(0.0); vt = Al.vector(0.0); zt = Al.vector(0.0)
        f.numpy()[:] = rng.standard_normal(n)
        u0 = rng.standard_normal(n)
        l1 = amg.level_array("l1", lev, n)
        for zero in (0, 1):
            u.numpy()[:] = 0.0 if zero else u0
            u.set_all_zeros_flag(zero)
            L.hypre_BoomerAMGRelax(Al.h, f.h, None, rtype, 0, C.c_double(amg.relax_weight(lev)), C.c_double(1.0),
                                   O.p(l1), u.h, vt.h, zt.h)
            ustart = np.zeros(n) if zero else u0
            if rtype == 18:
                uo = O.relax_l1jacobi(Ad, f.numpy(), l1, amg.relax_weight(lev), ustart, zero)
            else:
                uo, err = O.relax_gs2(Ad, f.numpy(), amg.relax_weight(lev), 1 if rtype == 11 else 2, ustart)
                assert err == 0
            np.testing.assert_allclose(uo, u.numpy(), rtol=1e-12, atol=1e-13)


def test_oracle_cheby_and_cycle_vs_reference(ref_off):
    H = ref_off
    L = H.lib()
    A, b, x, amg, pcg = O.build_case(H, "lap7_24_rlx16")
    prm = amg.params()
    rng = np.random.default_rng(3)
    Al = amg.A(0)
    Ad, _ = Al.blocks()
    n = Ad.shape[0]
    f = Al.vector(0.0); u = Al.vector(0.0)
    w = [Al.vector(0.0) for _ in range(4)]
    f.numpy()[:] = rng.standard_normal(n)
    u0 = rng.standard_normal(n)
    u.numpy()[:] = u0
    ds = amg.level_array("cheby_ds", 0, n)
    coefs = amg.level_array("cheby_coefs", 0, prm["cheby_order"] + 1)
    L.hypre_ParCSRRelax_Cheby_Solve(Al.h, f.h, O.p(ds), O.p(coefs), prm["cheby_order"], prm["cheby_scale"],
                                    prm["cheby_variant"], u.h, w[0].h, w[1].h, w[2].h, w[3].h)
    uo = O.relax_cheby(Ad, f.numpy(), ds, coefs, prm["cheby_order"], prm["cheby_scale"], u0)
    np.testing.assert_allclose(uo, u.numpy(), rtol=1e-12, atol=1e-13)
    # one full V-cycle through HYPRE_BoomerAMGSolve (tol 0, max_iter 1) vs orc_amg_cycle
    orc = O.OracleAMG(amg)
    u.numpy()[:] = u0
    u.set_all_zeros_flag(0)
    amg.Solve(A, f, u)
    uc = orc.cycle(f.numpy(), u0, 0)
    np.testing.assert_allclose(uc, u.numpy(), rtol=1e-10, atol=1e-12)
    orc.destroy()


def test_gselim_restatement():
    rng = np.random.default_rng(4)
    for n in (1, 2, 5, 22):
        M = rng.standard_normal((n, n)) + n * np.eye(n)
        rhs = rng.standard_normal(n)
        A = np.ascontiguousarray(M.copy()); x = rhs.copy()
        assert O.lib().orc_gselim(O.p(A), O.p(x), n) == 0
        np.testing.assert_allclose(x, np.linalg.solve(M, rhs), rtol=1e-10)
    A = np.zeros((1, 1)); x = np.ones(1)
    assert O.lib().orc_gselim(O.p(A), O.p(x), 1) == 1


@pytest.mark.parametrize("inner", [1, 2])
def test_gs2_block_partitioned_restatement(ref_off, inner):
    """N-rank two-stage GS: the inner stages use the strict lower triangle of each rank's diag block.
    One block must reproduce the single-rank routine; several blocks are checked against a matrix
    formulation z_{k+1} = D^{-1} L_blocks z_k, u += (-1)^k z_k."""
    import scipy.sparse as sp
    H = ref_off
    A = H.laplacian_27pt(9, 8, 7).blocks()[0]
    n = A.shape[0]
    ii, jj, aa = O.csr_arrays(A)
    rng = np.random.default_rng(3)
    f, u0 = rng.standard_normal(n), rng.standard_normal(n)
    L = O.lib()

    def run(starts):
        u = u0.copy(); vt = np.zeros(n)
        st = np.ascontiguousarray(starts, dtype=np.int32)
        err = L.orc_relax_gs2_blocks(n, O.p(ii), O.p(jj), O.p(aa), O.p(f), 1.0, inner, O.p(u), O.p(vt), len(st) - 1, O.p(st))
        assert err == 0
        return u

    u_one, err = O.relax_gs2(A, f, 1.0, inner, u0)
    assert np.array_equal(run([0, n]), u_one)
    starts = [0, 100, 230, 231, n]
    d = A.diagonal()
    Lb = sp.tril(A, k=-1).tocoo()
    blk = np.searchsorted(starts, np.arange(n), side="right") - 1
    keep = blk[Lb.row] == blk[Lb.col]
    Lb = sp.csr_matrix((Lb.data[keep], (Lb.row[keep], Lb.col[keep])), shape=(n, n))
    z = (f - A @ u0) / d
    u = u0 + z
    sign = -1.0
    for _ in range(inner):
        z = (Lb @ z) / d
        u = u + sign * z
        sign = -sign
    np.testing.assert_allclose(run(starts), u, rtol=1e-13, atol=1e-14)
    assert not np.allclose(run(starts), u_one, rtol=1e-6)      # the partition does matter


@pytest.mark.parametrize("name", ["lap7_24_rlx18_W", "lap7_24_rlx16_F", "lap27_20_rlx18_ns2", "lap7_24_rlx7_w08",
                                  "lap7_24_rlx16_var1", "lap7_24_rlx16_o4", "lap7_24_rlx16_o3_noscale",
                                  "lap7_24_rlx18_pc3", "lap7_24_rlx16_pc2", "lap7_24_rlx0_w08", "lap27_20_rlx0_w07_ns2",
                                  "lap7_24_rlx18_CF", "lap27_20_rlx0_CF_w07", "lap7_24_rlx7_CF_W", "lap27_20_rlx11_CF",
                                  "lap7_24_rlx16_CF"])
def test_oracle_pcg_option_cases_match_reference(ref_off, name):
    """cycle and smoother options without a golden (W / F cycle, two sweeps, weighted Jacobi, Chebyshev variant 1,
    orders 3 / 4, unscaled): the oracle against hypre_PCGSolve in-process on the same hierarchy"""
    H = ref_off
    A, b, x, amg, pcg = O.build_case(H, name)
    H.clear_errors()
    pcg.Solve(A, b, x)
    assert H.error_flag() == 0
    orc = O.OracleAMG(amg)
    xs, it, rel, conv = orc.pcg(b.numpy(), np.zeros(x.size))
    assert it == pcg.num_iterations and conv == 1
    np.testing.assert_allclose(rel, pcg.rel_norms(), rtol=1e-9)
    np.testing.assert_allclose(xs, x.numpy(), rtol=1e-9, atol=1e-12)
    orc.destroy()
