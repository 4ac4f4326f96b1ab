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


# ------------------------------------------------------------------ compressed thread-per-row storage
def _storage(L, dA):
    fmt, nd = C.c_int(), C.c_int()
    sb, cb = C.c_longlong(), C.c_longlong()
    capi.check(L.b200_parcsr_storage(dA, C.byref(fmt), C.byref(sb), C.byref(cb), C.byref(nd)))
    return fmt.value, sb.value, cb.value, nd.value


_COMP_CACHE = {}


def _random_matrices():
    """built once (csr arithmetic only: lil assignment of wide slices is very slow)"""
    if "r" not in _COMP_CACHE:
        rng = np.random.default_rng(14)
        # many distinct values, columns far apart (32-bit escape groups), empty rows, a long row
        # (explicit coordinates: sp.random draws without replacement from rows*cols, minutes at this shape)
        ri = rng.integers(0, 3000, 30000); cj = rng.integers(0, 400000, 30000)
        ok = (ri < 200) | (ri >= 260)                                   # empty rows
        cols = np.arange(0, 400000, 97)                                 # one long row
        ri = np.concatenate([ri[ok], np.full(len(cols), 7)]); cj = np.concatenate([cj[ok], cols])
        A = sp.csr_matrix((rng.standard_normal(len(ri)), (ri, cj)), shape=(3000, 400000))
        A.sum_duplicates(); A.sort_indices()
        B = sp.random(5000, 5000, density=0.004, random_state=6, format="csr")
        Cm = B.copy(); Cm.data = np.round(Cm.data * 200.0) / 7.0        # ~200 distinct values, random columns
        _COMP_CACHE["r"] = (A, B, Cm)
    return _COMP_CACHE["r"]


def _compression_matrices(H, hier):
    amg = hier[3]
    yield "lap7", H.laplacian(19, 17, 13).blocks()[0], 4            # 7 (delta, value) patterns -> pattern storage
    yield "lap27", H.laplacian_27pt(14, 12, 11).blocks()[0], 4
    yield "A1", amg.A(1).blocks()[0], None
    yield "P0", amg.P(0).blocks()[0], None
    A, B, Cm = _random_matrices()
    yield "wide_u16", A, 2                                          # <= 65536 distinct values
    yield "distinct_f64", B, 3                                      # > 65536 distinct values
    yield "random_u8", Cm, 1


# (mode, no_padding_markers, chunked, pat_regs): every storage format in both padding forms, the pattern kernel in
# both instantiations, both slice-to-block mappings
_FORMATS = [(-1, 0, 1, 2), (0, 0, 1, 2), (1, 0, 1, 2), (2, 0, 1, 2), (3, 0, 1, 2), (4, 0, 1, 2), (4, 0, 0, 9), (4, 0, 1, 3),
            (-1, 1, 1, 8), (1, 1, 0, 2), (2, 1, 1, 2), (3, 1, 1, 2), (1, 0, 2, 2)]


@pytest.mark.parametrize("layout", [1, 2])
@pytest.mark.parametrize("mode,zpad,chunked,pat_regs", _FORMATS)
def test_compressed_storage_is_lossless(H, ctx, hier, layout, mode, zpad, chunked, pat_regs):
    """16-bit column offsets + value dictionary (b200_parcsr_storage): same products, same order as
    the CSR row of the reference (seq_mv/csr_matvec.c:21-857) -> SpMV parity <= 1e-12 in every format."""
    L = capi.lib()
    rng = np.random.default_rng(15)
    capi.check(L.b200_set_layout_override(layout, -1))
    capi.check(L.b200_set_compression(mode, zpad))
    capi.check(L.b200_set_kernel_options(chunked, pat_regs))
    try:
        for nm, A, auto_fmt in _compression_matrices(H, hier):
            n, m = A.shape
            dA = capi.parcsr_from_scipy(ctx, A)
            fmt, stored, csr, nd = _storage(L, dA)
            if mode == 0:
                assert fmt == 0
            elif mode in (-1, 4):
                assert fmt in (0, 1, 2, 3, 4)
                if fmt and auto_fmt:
                    assert fmt == auto_fmt, (nm, fmt)
                if nm == "A1":
                    assert fmt == 4, fmt                       # first Galerkin product of a stencil: still patterns
                if fmt and nm in ("lap7", "lap27", "A1"):      # regular rows: far fewer bytes than CSR
                    assert stored < 0.2 * csr, (nm, stored, csr)
            else:
                assert fmt in (mode, 3) or (mode == 2 and fmt == 2), (nm, mode, fmt)
            x = rng.standard_normal(m); b = rng.standard_normal(n)
            dx, db, dy = dvec(ctx, x), dvec(ctx, b), capi.DeviceVector(ctx, n)
            for alpha, beta in [(1.0, 0.0), (-1.0, 1.0)]:
                capi.check(L.b200_parcsr_matvec(dA, alpha, dx.ptr, beta, db.ptr, dy.ptr))
                assert row_err(dy.get(), O.matvec(A, alpha, x, beta, b), A, x) < SPMV_TOL, (nm, layout, mode, zpad, fmt)
            # every format computes the CSR row's fma chain: bit-identical to the plain layout
            dx.set(x)
            capi.check(L.b200_parcsr_matvec(dA, 1.0, dx.ptr, 0.0, db.ptr, dy.ptr))
            key = (nm, layout)
            if key in _BITS:
                assert np.array_equal(dy.get(), _BITS[key]), (nm, layout, mode, zpad, fmt)
            else:
                _BITS[key] = dy.get()
            # NaN in a column a row does not touch must not leak through padding lanes
            xn = x.copy(); cols = np.setdiff1d(np.arange(m), A.indices)
            if len(cols):
                xn[cols[0]] = np.nan
                dx.set(xn)
                capi.check(L.b200_parcsr_matvec(dA, 1.0, dx.ptr, 0.0, db.ptr, dy.ptr))
                assert np.all(np.isfinite(dy.get())), nm
            for v in (dx, db, dy):
                v.free()
            L.b200_parcsr_destroy(dA)
    finally:
        L.b200_set_layout_override(-1, -1)
        L.b200_set_compression(-1, 1)
        L.b200_set_kernel_options(1, 8)


_BITS = {}      # SpMV result of the first format run on (matrix, layout): the others must reproduce it bit for bit


@pytest.mark.parametrize("mode,zpad,chunked,pat_regs", [(-1, 0, 1, 2), (4, 0, 0, 9), (1, 0, 1, 2), (1, 1, 1, 2), (2, 1, 0, 2), (3, 0, 1, 2), (3, 1, 1, 2)])
def test_smoothers_on_compressed_storage(H, ctx, hier, mode, zpad, chunked, pat_regs):
    """every fused epilogue on the compressed thread-per-row kernels against the oracle"""
    L = capi.lib()
    rng = np.random.default_rng(16)
    amg = hier[3]
    capi.check(L.b200_set_layout_override(1, -1))
    capi.check(L.b200_set_compression(mode, zpad))
    capi.check(L.b200_set_kernel_options(chunked, pat_regs))
    try:
        for l in (0, 1):
            A = amg.A(l).blocks()[0]
            n = A.shape[0]
            dA = capi.parcsr_from_scipy(ctx, A)
            assert _storage(L, dA)[0] != 0
            f = rng.standard_normal(n)
            l1 = amg.level_array("l1", l, n)
            ds = 1.0 / np.sqrt(np.abs(A.diagonal()))
            coefs = np.array([0.9, -0.31, 0.077])
            for zg in (0, 1):
                u0 = np.zeros(n) if zg else rng.standard_normal(n)
                df, dl1, dds, du = dvec(ctx, f), dvec(ctx, l1), dvec(ctx, ds), dvec(ctx, u0)
                w = [capi.DeviceVector(ctx, n) for _ in range(3)]
                capi.check(L.b200_relax_l1jacobi(dA, df.ptr, dl1.ptr, 0.9, du.ptr, w[0].ptr, zg))
                assert vec_err_max(du.get(), O.relax_l1jacobi(A, f, l1, 0.9, u0, zg)) < 1e-11
                du.set(u0)
                capi.check(L.b200_relax_cheby(dA, df.ptr, dds.ptr, capi.ptr(coefs), 2, 1, du.ptr, w[0].ptr, w[1].ptr, w[2].ptr, zg))
                assert vec_err_max(du.get(), O.relax_cheby(A, f, ds, coefs, 2, 1, u0)) < 1e-11
                du.set(u0)
                capi.check(L.b200_relax_2stage_gs(dA, df.ptr, 1.0, 1, du.ptr, w[0].ptr, w[1].ptr, zg))
                uref, err = O.relax_gs2(A, f, 1.0, 1, u0)
                assert err == 0 and vec_err_max(du.get(), uref) < 1e-11
                for v in [df, dl1, dds, du] + w:
                    v.free()
            L.b200_parcsr_destroy(dA)
    finally:
        L.b200_set_layout_override(-1, -1)
        L.b200_set_compression(-1, 1)
        L.b200_set_kernel_options(1, 8)


def vec_err_max(a, b):
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))
