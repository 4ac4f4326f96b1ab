"""Host-side tests of the HBM storage formats (hypre_b200/csrc/pack.cu): layout choice, sliced ELL /
SELL-32-256 packing and the lossless compressed form (16-bit column offsets + value dictionary).
b200_storage_decode_host packs a CSR block exactly as b200_parcsr_create does and multiplies it with a
vector FROM THE PACKED ARRAYS on the host, with the kernels' addressing and fma chain; no GPU needed.
The CSR row of the reference (seq_mv/csr_matvec.c:21-857) is the yardstick: every format must give the
bit-identical row sums of the plain CSR fma chain."""
import ctypes as C

import numpy as np
import pytest
import scipy.sparse as sp

from hypre_b200 import capi


def decode(A, x, layout, mode, zpad=0):
    L = capi.lib()
    A = sp.csr_matrix(A)
    n, m = A.shape
    ai = np.ascontiguousarray(A.indptr, dtype=np.int32)
    aj = np.ascontiguousarray(A.indices, dtype=np.int32)
    aa = np.ascontiguousarray(A.data, dtype=np.float64)
    y = np.full(n, np.nan)
    lay, fmt = C.c_int(), C.c_int()
    stored, nesc = C.c_longlong(), C.c_longlong()
    capi.check(L.b200_set_layout_override(layout, -1))
    capi.check(L.b200_set_compression(mode, zpad))
    try:
        capi.check(L.b200_storage_decode_host(n, m, capi.ptr(ai), capi.ptr(aj), capi.ptr(aa), capi.ptr(x), capi.ptr(y),
                                              C.byref(lay), C.byref(fmt), C.byref(stored), C.byref(nesc)))
    finally:
        L.b200_set_layout_override(-1, -1)
        L.b200_set_compression(-1, 1)
    return y, lay.value, fmt.value, stored.value, nesc.value


def stencil7(nx, ny, nz):
    """7-point Laplacian, diagonal first (hypre's generator order is irrelevant here)"""
    idx = np.arange(nx * ny * nz).reshape(nz, ny, nx)
    rows, cols, vals = [idx.ravel()], [idx.ravel()], [np.full(idx.size, 6.0)]
    for ax in range(3):
        for sh in (-1, 1):
            src = [slice(None)] * 3; dst = [slice(None)] * 3
            src[ax] = slice(1, None) if sh == -1 else slice(0, -1)
            dst[ax] = slice(0, -1) if sh == -1 else slice(1, None)
            rows.append(idx[tuple(src)].ravel()); cols.append(idx[tuple(dst)].ravel())
            vals.append(np.full(rows[-1].size, -1.0))
    return sp.csr_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(idx.size,) * 2)


def matrices():
    rng = np.random.default_rng(3)
    yield "stencil7", stencil7(40, 33, 9)                                     # 2 distinct values
    yield "stencil7_wide", stencil7(300, 300, 2)[:9000, :]                    # offsets +-90000: escape groups
    ri = rng.integers(0, 3000, 30000); cj = rng.integers(0, 400000, 30000)
    ok = (ri < 200) | (ri >= 260)
    cols = np.arange(0, 400000, 97)
    ri = np.concatenate([ri[ok], np.full(len(cols), 7)]); cj = np.concatenate([cj[ok], cols])
    A = sp.csr_matrix((rng.standard_normal(len(ri)), (ri, cj)), shape=(3000, 400000)); A.sum_duplicates()
    yield "wide_ragged", A                                                    # empty rows, one long row, escapes
    B = sp.random(5000, 5000, density=0.004, random_state=6, format="csr")
    yield "many_values", B                                                    # > 65536 distinct values
    Cm = B.copy(); Cm.data = np.round(Cm.data * 200.0) / 7.0
    yield "few_values", Cm
    yield "one_by_one", sp.csr_matrix(np.array([[4.0]]))
    yield "short_rows", sp.csr_matrix((np.ones(70), (np.arange(70), (np.arange(70) * 7) % 50)), shape=(70, 50))
    Z = sp.random(100, 90, density=0.05, random_state=8, format="csr"); Z.data[::5] = -0.0; Z.data[1::5] = 0.0
    yield "signed_zeros", Z


@pytest.mark.parametrize("name,A", list(matrices()))
def test_every_format_decodes_to_the_csr_row_sums(name, A):
    rng = np.random.default_rng(4)
    A = sp.csr_matrix(A)
    x = rng.standard_normal(A.shape[1])
    y_csr, lay, fmt, stored_csr, _ = decode(A, x, 0, 0)
    assert lay == 0 and fmt == 0
    # the plain CSR chain against numpy (different summation order: tolerance), then bit-exactness between formats
    ref = A.dot(x)
    scale = np.abs(A).dot(np.abs(x)) + 1e-300
    assert np.max(np.abs(y_csr - ref) / scale) < 1e-14
    seen = set()
    for layout in (1, 2):
        for mode in (0, -1, 1, 2, 3, 4):
            y, lay, fmt, stored, nesc = decode(A, x, layout, mode)
            assert lay == layout
            assert np.array_equal(y, y_csr), (name, layout, mode, fmt)
            if mode == 0:
                assert fmt == 0
            elif mode in (1, 2, 3):
                assert fmt in (mode, 3) if mode != 2 else fmt in (2, 3)
            seen.add(fmt)
            if name in ("stencil7_wide", "wide_ragged") and fmt in (1, 2, 3):
                assert nesc > 0
    if name == "stencil7":
        y, lay, fmt, stored, nesc = decode(A, x, 1, 1)
        assert fmt == 1 and nesc == 0 and stored < 0.4 * stored_csr
        # pattern storage: 7 (delta, value) pairs + their padding twins; interior slices store a header only
        y, lay, fmt, stored4, nesc = decode(A, x, 1, -1)
        assert fmt == 4 and stored4 < 0.5 * stored
    if name in ("wide_ragged", "many_values", "signed_zeros"):
        assert 4 not in seen            # empty rows / too many patterns: not eligible
    if name == "many_values":
        assert 1 not in seen and 2 not in seen and 3 in seen
    if name == "few_values":
        assert decode(A, x, 1, -1)[2] == 1


def test_padding_lanes_do_not_touch_other_columns():
    """NaN in a column no row references must not reach any row sum (padding lanes are skipped or gather a
    column of their own row)"""
    A = stencil7(20, 10, 5)[:, :].tolil(); A = sp.csr_matrix(A)
    A = sp.hstack([A, sp.csr_matrix((A.shape[0], 3))]).tocsr()      # three untouched columns
    x = np.ones(A.shape[1]); x[-2] = np.nan
    for layout in (1, 2):
        for mode in (0, -1, 2, 3, 4):
            y = decode(A, x, layout, mode)[0]
            assert np.all(np.isfinite(y)), (layout, mode)


@pytest.mark.parametrize("name,A", list(matrices()))
def test_storage_without_padding_markers(name, A):
    """storage without padding markers (b200_set_compression(mode, 1) before the upload): padding lanes are (own column) x (a
    dictionary entry 0.0) instead of a marker.  Same row sums (a sum of -0.0 may become +0.0, which
    array_equal does not see), one more dictionary entry at most, and a NaN in a column a row does not
    reference still cannot reach that row."""
    rng = np.random.default_rng(5)
    A = sp.csr_matrix(A)
    x = rng.standard_normal(A.shape[1])
    y_csr = decode(A, x, 0, 0)[0]
    lens = np.diff(A.indptr)
    for layout in (1, 2):
        for mode in (-1, 1, 2, 3, 4):
            y, lay, fmt, stored, nesc = decode(A, x, layout, mode, zpad=1)
            y0, _, fmt0, stored0, nesc0 = decode(A, x, layout, mode, zpad=0)
            assert lay == layout and np.array_equal(y, y_csr), (name, layout, mode, fmt)
            assert nesc >= nesc0 and stored >= stored0
            if nesc == nesc0 and fmt == fmt0:
                assert stored == stored0                 # same arrays, only their padding differs
            # poison one column, rows that do not reference it (and have entries) stay finite
            col = int(A.indices[len(A.indices) // 2]) if A.nnz else 0
            xp = x.copy(); xp[col] = np.nan
            yp = decode(A, xp, layout, mode, zpad=1)[0]
            touched = np.zeros(A.shape[0], dtype=bool)
            touched[np.repeat(np.arange(A.shape[0]), lens)[A.indices == col]] = True
            assert np.all(np.isnan(yp[touched]))
            assert np.all(np.isfinite(yp[~touched & (lens > 0)])), (name, layout, mode)


def test_no_marker_storage_keeps_the_u8_dictionary_within_256_codes():
    """255 distinct values + the padding zero still fit u8 codes; 256 distinct values move to u16"""
    rng = np.random.default_rng(9)
    n = 600
    for distinct, want in ((255, 1), (256, 2)):
        vals = np.arange(1, distinct + 1, dtype=np.float64)
        rows = np.repeat(np.arange(n), 3); cols = (rows * 7 + np.tile(np.arange(3), n) * 11) % n
        keep = rng.random(rows.size) < 0.8                       # ragged: padding lanes exist
        keep[:distinct] = True
        data = np.resize(vals, rows.size)
        A = sp.csr_matrix((data[keep], (rows[keep], cols[keep])), shape=(n, n)); A.sum_duplicates()
        assert len(np.unique(A.data)) >= distinct or distinct == 256
        x = rng.standard_normal(n)
        y, lay, fmt, _, _ = decode(A, x, 1, 1 if want == 1 else 2, zpad=1)
        assert np.array_equal(y, decode(A, x, 0, 0)[0])
        if len(np.unique(A.data)) == distinct:
            assert fmt == want


def test_automatic_choice_keeps_small_blocks_in_csr_and_compresses_large_stencils():
    x = np.ones(40 * 33 * 9)
    assert decode(stencil7(40, 33, 9), x, -1, -1)[1] == 0             # < 131072 rows: CSR-vector
    A = stencil7(64, 64, 33)                                          # 135168 rows
    y, lay, fmt, stored, nesc = decode(A, np.ones(A.shape[0]), -1, -1)
    assert lay == 1 and fmt == 4 and nesc == 0
    assert stored < 1.0 * A.nnz                                       # pattern storage: well under a byte per entry
    assert np.array_equal(y, decode(A, np.ones(A.shape[0]), 0, 0)[0])


def test_pattern_storage_limits_and_shapes():
    """CFMT_PAT8 eligibility: <= 254 (delta, value) pairs including the padding twins, no empty rows; works for
    rectangular blocks whose columns follow the rows, and with the sorted-window layout (row permutation)."""
    rng = np.random.default_rng(11)
    n = 4000
    # banded matrix with `nd` distinct values on a fixed 3-point pattern: 3 * nd patterns (no padding: equal rows)
    for nd, want in ((84, 4), (85, 1)):          # 252 patterns + dead-lane twin fit; 255 + twin do not
        vals = np.arange(1, nd + 1, dtype=np.float64)
        rows = np.repeat(np.arange(n), 3)
        cols = (rows + np.tile(np.array([0, 5, 11]), n)) % (n + 11)
        data = vals[(rows * 7 + cols) % nd]
        A = sp.csr_matrix((data, (rows, cols)), shape=(n, n + 11))
        # make every (delta, value) combination appear
        x = rng.standard_normal(n + 11)
        y, lay, fmt, stored, _ = decode(A, x, 1, -1)
        assert np.array_equal(y, decode(A, x, 0, 0)[0])
        pats = len(set(zip((cols - rows).tolist(), data.tolist())))
        if pats + 1 <= 254:
            assert fmt == 4, (nd, pats, fmt)
        elif pats > 254:
            assert fmt != 4, (nd, pats, fmt)
    # ragged rows (boundary of a stencil), both thread-per-row layouts
    A = stencil7(37, 11, 5)
    x = rng.standard_normal(A.shape[1])
    for layout in (1, 2):
        y, lay, fmt, stored, _ = decode(A, x, layout, 4)
        assert lay == layout and fmt == 4 and np.array_equal(y, decode(A, x, 0, 0)[0])
    # an empty row disqualifies the block (its padding would have no column of its own)
    B = A.tolil(); B[5, :] = 0; B = sp.csr_matrix(B); B.eliminate_zeros()
    y, lay, fmt, _, _ = decode(B, x, 1, 4)
    assert fmt != 4 and np.array_equal(y, decode(B, x, 0, 0)[0])


def test_pattern_storage_on_a_real_hierarchy():
    """levels 0 and 1 of the HMIS hierarchy of the 7-pt Laplacian qualify (7 and ~150 patterns); the interior
    slices of level 0 store a header only (with 64 points per line every slice touches the boundary here: one byte per entry)"""
    from hypre_b200 import hypre as H
    H.set_device_path(False)
    try:
        A = H.laplacian(64, 64, 32)
        b, x = A.vector(1.0), A.vector(0.0)
        amg = H.BoomerAMG(relax_type=18)
        amg.Setup(A, b, x)
        rng = np.random.default_rng(12)
        for l, max_bpe in ((0, 1.05), (1, 1.4)):
            M = amg.A(l).blocks()[0]
            xv = rng.standard_normal(M.shape[1])
            y, lay, fmt, stored, _ = decode(M, xv, 1, -1)
            assert fmt == 4, (l, fmt)
            assert stored < max_bpe * M.nnz, (l, stored / M.nnz)
            assert np.array_equal(y, decode(M, xv, 0, 0)[0])
        P0 = amg.P(0).blocks()[0]
        assert decode(P0, rng.standard_normal(P0.shape[1]), 2, -1)[2] in (1, 2, 3)    # interpolation: no patterns
        amg.destroy()
    finally:
        H.set_device_path(True)
