#!/usr/bin/env python
"""Profiling target: the level-0 row kernels (SpMV, l1-Jacobi sweep, Chebyshev step, two-stage GS)
on one stencil matrix, without an AMG setup.  Used under ncu (--set full) and stand-alone, where it
prints CUDA-event GB/s per operator against the algorithmic bytes of DESIGN.md."""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hypre_b200 import capi, hypre as H   # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gen", default="laplacian", choices=["laplacian", "laplacian_27pt"])
    ap.add_argument("--n", type=int, default=256)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--layout", type=int, default=-1)
    ap.add_argument("--lanes", type=int, default=-1)
    ap.add_argument("--compress", type=int, default=-1, help="b200_set_compression mode (-1 auto, 0 off, 1..3 value formats, 4 patterns)")
    ap.add_argument("--zpad", type=int, default=-1, help="1: value formats packed without padding markers")
    a = ap.parse_args()
    L = capi.lib()
    capi.check(L.b200_set_layout_override(a.layout, a.lanes))
    capi.check(L.b200_set_compression(a.compress, a.zpad))
    print(json.dumps(measure(a.gen, a.n, a.reps)))


def measure(gen, size, reps):
    """level-0 operators of one generated stencil matrix: ms, CSR-equivalent GB/s, bytes (bench.py's aux leg too)"""
    a = argparse.Namespace(gen=gen, n=size, reps=reps)
    L = capi.lib()
    HL = H.lib()
    ctx = C.c_void_p(HL.hypre_b200_context())
    A = getattr(H, a.gen)(a.n, a.n, a.n)
    dA = A.device_mirror()
    n = A.global_rows
    nnz = C.c_longlong(); layout = C.c_int(); lanes = C.c_int()
    capi.check(L.b200_parcsr_info(dA, None, None, C.byref(nnz), C.byref(layout), C.byref(lanes)))
    nnz = nnz.value
    Ad = None
    x = np.sin(np.arange(n)) * 1e-3 + 1.0
    v = [capi.DeviceVector(ctx, n, x) for _ in range(7)]
    coefs = np.array([0.9, -0.3, 0.05])
    ms = C.c_float()
    spmv = 12.0 * nnz + 4.0 * n

    def timed(fn, bytes_):
        for _ in range(3):
            fn()
        capi.check(L.b200_timer_start(ctx))
        for _ in range(a.reps):
            fn()
        capi.check(L.b200_timer_stop(ctx, C.byref(ms)))
        t = ms.value / a.reps
        return {"ms": t, "GB/s": bytes_ / t / 1e6, "MB": bytes_ / 1e6}

    vfmt = C.c_int(); st_b = C.c_longlong(); csr_b = C.c_longlong(); ndict = C.c_int()
    capi.check(L.b200_parcsr_storage(dA, C.byref(vfmt), C.byref(st_b), C.byref(csr_b), C.byref(ndict)))
    out = {"gen": a.gen, "n": a.n, "rows": n, "nnz": nnz, "layout": layout.value, "lanes": lanes.value,
           "value_format": vfmt.value, "matrix_stored_MB": st_b.value / 1e6,
           "matrix_csr_MB": csr_b.value / 1e6, "note": "GB/s count the CSR (algorithmic) bytes"}
    out["spmv"] = timed(lambda: capi.check(L.b200_parcsr_matvec(dA, 1.0, v[0].ptr, 0.0, None, v[1].ptr)),
                        spmv + 16.0 * n)
    out["residual"] = timed(lambda: capi.check(L.b200_parcsr_matvec(dA, -1.0, v[0].ptr, 1.0, v[2].ptr, v[1].ptr)),
                            spmv + 24.0 * n)
    out["l1jacobi"] = timed(lambda: capi.check(L.b200_relax_l1jacobi(dA, v[2].ptr, v[3].ptr, 1.0, v[0].ptr, v[1].ptr, 0)),
                            spmv + 32.0 * n + 16.0 * n)   # + the copy-back of the stand-alone entry point
    out["cheby2"] = timed(lambda: capi.check(L.b200_relax_cheby(dA, v[2].ptr, v[3].ptr, capi.ptr(coefs), 2, 1,
                                                                v[0].ptr, v[4].ptr, v[5].ptr, v[6].ptr, 0)),
                          2 * (spmv + 40.0 * n))
    out["gs11"] = timed(lambda: capi.check(L.b200_relax_2stage_gs(dA, v[2].ptr, 1.0, 1, v[0].ptr, v[4].ptr, v[5].ptr, 0)),
                        spmv + 32.0 * n + 6.0 * nnz + 44.0 * n)
    out["axpy"] = timed(lambda: capi.check(L.b200_vec_axpy(ctx, n, 0.5, v[0].ptr, v[1].ptr)), 24.0 * n)
    r = C.c_double()
    out["dot"] = timed(lambda: capi.check(L.b200_vec_dot(ctx, n, v[0].ptr, v[1].ptr, C.byref(r))), 16.0 * n)
    for vec in v:
        vec.free()
    A.destroy()
    return out


if __name__ == "__main__":
    main()
