"""b200_vec_dot_noalloc: the inner product whose reduction scratch is a file-scope static device array (no
cudaMalloc anywhere on its path; include/b200amg.h) -- no counterpart
in the reference.  Collected last (test_zz_*).

Checked through the C-ABI against the existing entry point b200_vec_dot on the same device vectors, within the
tolerance of tests/test_gpu_parity.py::test_blas1 (1e-12 relative to sum |x_i y_i|), on two calls in a row (the
shared scratch keeps the first call's contents: the second call, on other data and another size, must not see
them), plus the edge sizes and the rejection beyond the stated maximum."""
import ctypes as C

import numpy as np
import pytest

from hypre_b200 import capi

pytestmark = [pytest.mark.gpu]

MAX_N = 67108864          # B200_NOALLOC_DOT_MAX_N (include/b200amg.h)


def _both(L, ctx, x, y, dout):
    n = len(x)
    dx, dy = capi.DeviceVector(ctx, n, x), capi.DeviceVector(ctx, n, y)
    try:
        capi.check(L.b200_vec_dot_noalloc(ctx, n, dx.ptr, dy.ptr, dout.ptr))
        got = float(dout.get()[0])
        r = C.c_double()
        capi.check(L.b200_vec_dot(ctx, n, dx.ptr, dy.ptr, C.byref(r)))
    finally:
        dx.free(); dy.free()
    return got, r.value


def test_dot_without_allocation_matches_the_existing_entry_point(ctx):
    L = capi.lib()
    rng = np.random.default_rng(77)
    dout = capi.DeviceVector(ctx, 1, np.array([np.nan]))
    try:
        # two calls in a row on different data and sizes (the second grid is smaller than the first: stale partials
        # of the first call sit behind it in the scratch), then again the first: same bits as before
        cases = [rng.standard_normal((2, 1234567)), rng.standard_normal((2, 4097)) * 1e3, None]
        first = None
        for k, xy in enumerate(cases):
            x, y = cases[0] if xy is None else xy
            got, ref = _both(L, ctx, x, y, dout)
            scale = max(1.0, float(np.dot(np.abs(x), np.abs(y))))
            assert abs(got - ref) <= 1e-12 * scale, (k, got, ref)
            assert abs(got - float(np.dot(x, y))) <= 1e-12 * scale, (k, got)
            if k == 0:
                first = got
            if xy is None:
                assert got == first          # deterministic, untouched by the call in between
        # edge sizes: empty, one element, one block exactly, one past a block
        for n in (0, 1, 4096, 4097):
            x, y = rng.standard_normal(n), rng.standard_normal(n)
            got, ref = _both(L, ctx, x, y, dout) if n else (None, None)
            if n == 0:
                capi.check(L.b200_vec_dot_noalloc(ctx, 0, None, None, dout.ptr))
                assert dout.get()[0] == 0.0
            else:
                assert abs(got - ref) <= 1e-12 * max(1.0, float(np.dot(np.abs(x), np.abs(y)))), n
        # beyond the static scratch: a clear error, nothing launched, the output word untouched
        dout.set(np.array([42.0]))
        st = L.b200_vec_dot_noalloc(ctx, MAX_N + 1, dout.ptr, dout.ptr, dout.ptr)
        assert st != 0 and b"B200_NOALLOC_DOT_MAX_N" in L.b200_last_error()
        assert dout.get()[0] == 42.0
    finally:
        dout.free()
