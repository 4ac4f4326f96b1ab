#!/usr/bin/env python
"""Storage formats and kernel options of the thread-per-row row kernels, level by level on a real hierarchy:
for A_0 .. A_3 of `ij -laplacian -n N N N` (HMIS, ext+i) every configuration below is uploaded and the level's
operators (SpMV, residual, l1-Jacobi sweep, Chebyshev order 2) are timed with CUDA events; results must be
bit-identical between configurations.  One JSON line per (level, configuration).

    python scripts/fmt_probe.py --n 256 --levels 0,1,2 > gpurun_out/fmt_probe.jsonl
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hypre_b200 import capi, hypre as H   # noqa: E402

# name: (compression mode, no_padding_markers, [(chunked, pat_regs), ...] launch options timed on that upload)
CONFIGS = {
    # pat_regs: bits 1-3 = waves of resident blocks in the grid of the pattern kernel (8 = four waves, the default)
    # pat_regs: bits 1-3 = waves of resident blocks in the grid of the pattern kernel (8 = four waves, the default)
    "pat8": (-1, 1, [(1, 8), (0, 8), (1, 2)]),
    "val_markers": (1, 0, [(1, 0), (0, 0)]),
    "val_zpad": (1, 1, [(1, 0), (0, 0)]),
    "plain": (0, 0, [(1, 0), (0, 0)]),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gen", default="laplacian")
    ap.add_argument("--n", type=int, default=256)
    ap.add_argument("--levels", default="0,1,2")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--configs", default="pat8,val_markers,val_zpad")
    a = ap.parse_args()
    L = capi.lib()
    HL = H.lib()
    ctx = C.c_void_p(HL.hypre_b200_context())
    H.set_device_path(False)
    A = getattr(H, a.gen)(a.n, a.n, a.n)
    b, x = A.vector(1.0), A.vector(0.0)
    amg = H.BoomerAMG(relax_type=16)
    t0 = time.perf_counter()
    amg.Setup(A, b, x)
    sys.stderr.write("setup %.1f s, %d levels\n" % (time.perf_counter() - t0, amg.num_levels))
    ms = C.c_float()
    coefs = np.array([0.9, -0.3, 0.05])
    for lev in [int(v) for v in a.levels.split(",")]:
        if lev >= amg.num_levels:
            continue
        M = amg.A(lev).blocks()[0]
        n, nnz = M.shape[0], M.nnz
        rng = np.random.default_rng(lev)
        xv = np.sin(np.arange(n)) * 1e-3 + 1.0
        fv = np.cos(np.arange(n) * 0.37)
        vx, vf, vl1 = (capi.DeviceVector(ctx, n, h) for h in (xv, fv, np.full(n, 6.5)))
        w = [capi.DeviceVector(ctx, n, np.zeros(n)) for _ in range(5)]
        ops = {
            "spmv": (lambda dA: capi.check(L.b200_parcsr_matvec(dA, 1.0, vx.ptr, 0.0, None, w[0].ptr)), 16.0),
            "residual": (lambda dA: capi.check(L.b200_parcsr_matvec(dA, -1.0, vx.ptr, 1.0, vf.ptr, w[0].ptr)), 24.0),
            "l1jacobi": (lambda dA: capi.check(L.b200_relax_l1jacobi(dA, vf.ptr, vl1.ptr, 1.0, w[1].ptr, w[0].ptr, 0)), 48.0),
            "cheby2": (lambda dA: capi.check(L.b200_relax_cheby(dA, vf.ptr, vl1.ptr, capi.ptr(coefs), 2, 1,
                                                               w[1].ptr, w[2].ptr, w[3].ptr, w[4].ptr, 0)), 80.0),
        }
        first = {}
        for cname in a.configs.split(","):
            mode, zpad, opts = CONFIGS[cname]
            capi.check(L.b200_set_compression(mode, zpad))
            t0 = time.perf_counter()
            dA = capi.parcsr_from_scipy(ctx, M)
            upload_s = time.perf_counter() - t0
            vfmt = C.c_int(); st_b = C.c_longlong(); csr_b = C.c_longlong(); ndict = C.c_int()
            capi.check(L.b200_parcsr_storage(dA, C.byref(vfmt), C.byref(st_b), C.byref(csr_b), C.byref(ndict)))
            layout = C.c_int(); lanes = C.c_int()
            capi.check(L.b200_parcsr_info(dA, None, None, None, C.byref(layout), C.byref(lanes)))
            for chunked, pat_regs in opts:
                if pat_regs and vfmt.value != 4:
                    continue
                capi.check(L.b200_set_kernel_options(chunked, pat_regs))
                out = {"level": lev, "rows": n, "nnz": nnz, "config": cname, "chunked": chunked, "pat_regs": pat_regs,
                       "layout": layout.value, "value_format": vfmt.value, "n_dictionary": ndict.value,
                       "stored_MB": st_b.value / 1e6, "csr_MB": csr_b.value / 1e6, "upload_s": upload_s}
                for name, (fn, vec_bytes) in ops.items():
                    w[1].set(xv)
                    fn(dA)
                    got = (w[1] if name in ("l1jacobi", "cheby2") else w[0]).get()
                    if name in first:
                        out[name + "_identical"] = bool(np.array_equal(got, first[name]))
                    else:
                        first[name] = got
                    for _ in range(3):
                        fn(dA)
                    capi.check(L.b200_timer_start(ctx))
                    for _ in range(a.reps):
                        fn(dA)
                    capi.check(L.b200_timer_stop(ctx, C.byref(ms)))
                    t = ms.value / a.reps
                    passes = 2 if name == "cheby2" else 1
                    stored = passes * st_b.value + vec_bytes * n
                    out[name] = {"us": 1e3 * t, "stored_GBs": stored / t / 1e6, "csr_GBs": (passes * csr_b.value + vec_bytes * n) / t / 1e6}
                print(json.dumps(out), flush=True)
            capi.check(L.b200_parcsr_destroy(dA))
        for v in [vx, vf, vl1] + w:
            v.free()
    capi.check(L.b200_set_compression(-1, 1))
    capi.check(L.b200_set_kernel_options(1, 8))


if __name__ == "__main__":
    main()
