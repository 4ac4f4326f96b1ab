"""GPU parity tests (B200): every device operator against the plain-C oracle / the compiled
reference on the same inputs, through the C ABI (include/b200amg.h) and through the interposed
HYPRE API.  Tolerances are the north star's: SpMV <= 1e-12 relative per row, per-iteration
relative residual norms <= 1e-9, identical PCG iteration counts."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest
import scipy.sparse as sp

import oracle_util as O
from hypre_b200 import IJ_B200, IJ_REF, capi

pytestmark = pytest.mark.gpu

SPMV_TOL = 1e-12
NORM_TOL = 1e-9


def row_err(y, yref, A=None, x=None):
    """relative error per row, scaled by sum_j |a_ij x_j| (the conditioning of the row sum)"""
    if A is not None:
        scale = np.abs(A).dot(np.abs(x)) + np.abs(yref)
    else:
        scale = np.abs(yref)
    scale = np.maximum(scale, 1e-300)
    return float(np.max(np.abs(y - yref) / scale))


def dvec(ctx, a):
    return capi.DeviceVector(ctx, len(a), a)


def stencil_matrices(H):
    yield "lap7", H.laplacian(19, 17, 13)
    yield "lap27", H.laplacian_27pt(14, 12, 11)
    yield "rot", H.rotate_7pt(61, 47, 45.0, 0.001)


@pytest.fixture(scope="module")
def hier(H):
    """a real hierarchy (coarse levels have long ragged rows)"""
    H.set_device_path(True)
    return O.build_case(H, "lap7_32_rlx18")


@pytest.mark.parametrize("override", [(-1, -1), (1, -1), (2, -1), (0, 1), (0, 2), (0, 4), (0, 8), (0, 16), (0, 32)])
def test_spmv_matches_oracle_all_layouts(H, ctx, hier, override):
    L = capi.lib()
    rng = np.random.default_rng(10)
    mats = [(nm, A.blocks()[0]) for nm, A in stencil_matrices(H)]
    amg = hier[3]
    mats += [("A%d" % l, amg.A(l).blocks()[0]) for l in range(1, amg.num_levels)]
    mats += [("P%d" % l, amg.P(l).blocks()[0]) for l in range(0, 3)]
    capi.check(L.b200_set_layout_override(*override))
    try:
        for nm, A in mats:
            n, m = A.shape
            dA = capi.parcsr_from_scipy(ctx, A)
            x = rng.standard_normal(m); b = rng.standard_normal(n)
            dx, db, dy = dvec(ctx, x), dvec(ctx, b), capi.DeviceVector(ctx, n)
            for alpha, beta in [(1.0, 0.0), (-1.0, 1.0), (0.3, -2.0), (0.0, 1.5)]:
                capi.check(L.b200_parcsr_matvec(dA, alpha, dx.ptr, beta, db.ptr, dy.ptr))
                yref = O.matvec(A, alpha, x, beta, b)
                assert row_err(dy.get(), yref, A, x) < SPMV_TOL, (nm, override, alpha, beta)
            # y aliasing b (in-place residual form)
            capi.check(L.b200_vec_copy(ctx, n, db.ptr, dy.ptr))
            capi.check(L.b200_parcsr_matvec(dA, -1.0, dx.ptr, 1.0, dy.ptr, dy.ptr))
            assert row_err(dy.get(), O.matvec(A, -1.0, x, 1.0, b), A, x) < SPMV_TOL
            for v in (dx, db, dy):
                v.free()
            L.b200_parcsr_destroy(dA)
    finally:
        L.b200_set_layout_override(-1, -1)


def test_matvecT_matches_oracle(H, ctx, hier):
    L = capi.lib()
    rng = np.random.default_rng(11)
    amg = hier[3]
    for l in range(amg.num_levels - 1):
        P = amg.P(l).blocks()[0]
        n, m = P.shape
        dP = capi.parcsr_from_scipy(ctx, P)
        x = rng.standard_normal(n); y0 = rng.standard_normal(m)
        dx, dy = dvec(ctx, x), dvec(ctx, y0)
        for alpha, beta in [(1.0, 0.0), (1.0, 1.0), (-0.5, 2.0)]:
            dy.set(y0)
            capi.check(L.b200_parcsr_matvecT(dP, alpha, dx.ptr, beta, dy.ptr))
            yref = O.matvecT(P, alpha, x, beta, y0)
            assert row_err(dy.get(), yref, P.T.tocsr(), x) < SPMV_TOL
        dx.free(); dy.free(); L.b200_parcsr_destroy(dP)


def test_edge_cases_empty_single_ragged(ctx):
    L = capi.lib()
    # empty matrix
    dA = capi.parcsr_from_scipy(ctx, sp.csr_matrix((0, 0)))
    capi.check(L.b200_parcsr_matvec(dA, 1.0, None, 0.0, None, capi.DeviceVector(ctx, 0).ptr))
    L.b200_parcsr_destroy(dA)
    # 1 x 1
    A = sp.csr_matrix(np.array([[4.0]]))
    dA = capi.parcsr_from_scipy(ctx, A)
    dx, dy = dvec(ctx, np.array([2.0])), capi.DeviceVector(ctx, 1)
    capi.check(L.b200_parcsr_matvec(dA, 1.0, dx.ptr, 0.0, None, dy.ptr))
    assert dy.get()[0] == 8.0
    L.b200_parcsr_destroy(dA)
    # ragged rows: empty rows, one very long row, rectangular, n not a multiple of 32
    rng = np.random.default_rng(12)
    n, m = 1031, 777
    A = sp.random(n, m, density=0.01, random_state=3, format="lil")
    A[5, :] = rng.standard_normal(m)       # dense row
    A[100:140, :] = 0.0                    # empty rows
    A = A.tocsr(); A.eliminate_zeros()
    x = rng.standard_normal(m)
    for override in [(-1, -1), (1, -1), (2, -1), (0, 1), (0, 2), (0, 32)]:
        capi.check(L.b200_set_layout_override(*override))
        dA = capi.parcsr_from_scipy(ctx, A)
        dx, dy = dvec(ctx, x), capi.DeviceVector(ctx, n)
        capi.check(L.b200_parcsr_matvec(dA, 1.0, dx.ptr, 0.0, None, dy.ptr))
        assert row_err(dy.get(), O.matvec(A, 1.0, x, 0.0, None), A, x) < SPMV_TOL
        L.b200_parcsr_destroy(dA)
    L.b200_set_layout_override(-1, -1)


def test_blas1(ctx):
    L = capi.lib()
    rng = np.random.default_rng(13)
    for n in (0, 1, 31, 1000, 123457):
        x = rng.standard_normal(n); y = rng.standard_normal(n)
        dx, dy = dvec(ctx, x), dvec(ctx, y)
        capi.check(L.b200_vec_axpy(ctx, n, 0.75, dx.ptr, dy.ptr))
        np.testing.assert_allclose(dy.get(), y + 0.75 * x, rtol=1e-15, atol=1e-15)
        capi.check(L.b200_vec_scale(ctx, n, -2.0, dy.ptr))
        r = C.c_double()
        capi.check(L.b200_vec_dot(ctx, n, dx.ptr, dy.ptr, C.byref(r)))
        ref = float(np.dot(x, -2.0 * (y + 0.75 * x)))
        assert abs(r.value - ref) <= 1e-12 * max(1.0, float(np.dot(np.abs(x), np.abs(2.0 * (y + 0.75 * x)))))
        capi.check(L.b200_vec_fill(ctx, n, 3.5, dy.ptr))
        assert np.all(dy.get() == 3.5)
        capi.check(L.b200_vec_copy(ctx, n, dx.ptr, dy.ptr))
        assert np.array_equal(dy.get(), x)
