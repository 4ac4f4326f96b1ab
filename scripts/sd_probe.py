#!/usr/bin/env python
"""Round-2 candidate check: loop variants of the compressed sliced-ELL row kernel (4 = value dictionary in
shared memory, 5 = predicate-free loop on storage without padding markers; both opt-in) against the default
loop (variant 0) on the SAME device matrix: bit-identical results and CUDA-event times of SpMV / residual /
Chebyshev step / l1-Jacobi sweep.  The variant is chosen at launch time (b200_set_compression(-1, v)), so one
upload serves all; with 5 in --variants the matrix is packed without padding markers (variants 0 and 4 run on
that storage, too).  Round-1 result for variant 4: profiles/r01_sellc_variant4_probe.json."""
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
    ap.add_argument("--gen", default="laplacian")
    ap.add_argument("--n", type=int, default=256)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--variants", default="0,4")
    a = ap.parse_args()
    variants = [int(v) for v in a.variants.split(",")]
    assert variants[0] == 0
    L = capi.lib()
    HL = H.lib()
    ctx = C.c_void_p(HL.hypre_b200_context())
    A = getattr(H, a.gen)(a.n, a.n, a.n)
    if 5 in variants:
        capi.check(L.b200_set_compression(-1, 5))     # pack without padding markers
    dA = A.device_mirror()
    n = A.global_rows
    nnz = C.c_longlong(); layout = C.c_int(); lanes = C.c_int()
    capi.check(L.b200_parcsr_info(dA, None, None, C.byref(nnz), C.byref(layout), C.byref(lanes)))
    vfmt = C.c_int(); st_b = C.c_longlong(); csr_b = C.c_longlong(); ndict = C.c_int()
    capi.check(L.b200_parcsr_storage(dA, C.byref(vfmt), C.byref(st_b), C.byref(csr_b), C.byref(ndict)))
    out = {"gen": a.gen, "n": a.n, "rows": n, "nnz": nnz.value, "layout": layout.value,
           "value_format": vfmt.value, "n_dictionary": ndict.value, "stored_MB": st_b.value / 1e6,
           "csr_MB": csr_b.value / 1e6}
    x = np.sin(np.arange(n)) * 1e-3 + 1.0
    f = np.cos(np.arange(n) * 0.37)
    vx, vf, vl1 = (capi.DeviceVector(ctx, n, h) for h in (x, f, np.full(n, 6.5)))
    w = [capi.DeviceVector(ctx, n, np.zeros(n)) for _ in range(5)]
    coefs = np.array([0.9, -0.3, 0.05])
    ms = C.c_float()

    def ops():
        return {
            "spmv": lambda: capi.check(L.b200_parcsr_matvec(dA, 1.0, vx.ptr, 0.0, None, w[0].ptr)),
            "residual": lambda: capi.check(L.b200_parcsr_matvec(dA, -1.0, vx.ptr, 1.0, vf.ptr, w[0].ptr)),
            "l1jacobi": lambda: capi.check(L.b200_relax_l1jacobi(dA, vf.ptr, vl1.ptr, 1.0, w[1].ptr, w[0].ptr, 0)),
            "cheby2": lambda: capi.check(L.b200_relax_cheby(dA, vf.ptr, vl1.ptr, capi.ptr(coefs), 2, 1,
                                                            w[1].ptr, w[2].ptr, w[3].ptr, w[4].ptr, 0)),
        }

    results = {}
    for variant in variants:
        capi.check(L.b200_set_compression(-1, variant))
        res = {}
        for name, fn in ops().items():
            w[1].set(x)                      # the smoothers update u in place: same start for both variants
            fn()
            got = (w[1] if name in ("l1jacobi", "cheby2") else w[0]).get()
            for _ in range(3):
                fn()
            capi.check(L.b200_timer_start(ctx))
            for _ in range(a.reps):
                fn()
            capi.check(L.b200_timer_stop(ctx, C.byref(ms)))
            res[name] = (ms.value / a.reps, got)
        results[variant] = res
    out["padding_markers"] = 5 not in variants
    for name in results[0]:
        t0, y0 = results[0][name]
        out[name] = {"ms_variant0": t0}
        for v in variants[1:]:
            tv, yv = results[v][name]
            out[name].update({"ms_variant%d" % v: tv, "speedup_variant%d" % v: t0 / tv,
                              "bit_identical_variant%d" % v: bool(np.array_equal(y0, yv) and np.isfinite(yv).all()),
                              "max_abs_diff_variant%d" % v: float(np.max(np.abs(y0 - yv)))})
    print(json.dumps(out))


if __name__ == "__main__":
    main()
