#!/usr/bin/env python
"""Auxiliary timing of the device GMRES(k) / BiCGSTAB solves next to PCG on one GPU (not the bench.py contract
line: that stays AMG-PCG).  Same hierarchy, same right-hand side, device-resident vectors, CUDA events on the
library's stream; one JSON line per solver.

    python scripts/bench_krylov.py [--n 128] [--rlx 16] [--steps 3] [--k 5]
"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=128)
    ap.add_argument("--rlx", type=int, default=16)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--k", type=int, default=5)
    args = ap.parse_args()
    from hypre_b200 import capi, hypre as H
    L, HL = capi.lib(), H.lib()
    A = H.laplacian(args.n, args.n, args.n)
    b, x = A.vector(1.0), A.vector(0.0)
    amg = H.BoomerAMG(relax_type=args.rlx)
    amg.Setup(A, b, x)
    ctx = C.c_void_p(HL.hypre_b200_context())
    if not ctx:
        raise capi.B200Error("no CUDA device: " + L.b200_last_error().decode())
    damg = amg.device_mirror()
    dA = C.c_void_p()
    capi.check(L.b200_amg_level_matrix(damg, 0, C.byref(dA)))
    n = A.global_rows
    db, dx = capi.DeviceVector(ctx, n, b.numpy()), capi.DeviceVector(ctx, n)
    norms, rel = np.zeros(1001), np.zeros(1001)
    ms = C.c_float()

    def solver(kind):
        if kind == "pcg":
            prm, res = capi.PcgParams(1e-8, 0.0, 1000, 1, 0, 0), capi.PcgResult()
            return res, lambda: capi.check(L.b200_pcg_solve(ctx, dA, damg, C.byref(prm), db.ptr, dx.ptr, C.byref(res),
                                                            capi.ptr(norms), capi.ptr(rel)))
        if kind in ("gmres", "cogmres", "flexgmres"):
            prm = capi.GmresParams(1e-8, 0.0, 1000, 0, args.k, 0, 0, 1 if kind == "cogmres" else 0, 1 if kind == "flexgmres" else 0)
            res = capi.KrylovResult()
            return res, lambda: capi.check(L.b200_gmres_solve(ctx, dA, damg, C.byref(prm), db.ptr, dx.ptr, C.byref(res),
                                                              capi.ptr(norms)))
        if kind == "lgmres":
            prm, res = capi.LgmresParams(1e-8, 0.0, 1000, 0, args.k, 2, 1, 0), capi.KrylovResult()
            return res, lambda: capi.check(L.b200_lgmres_solve(ctx, dA, damg, C.byref(prm), db.ptr, dx.ptr, C.byref(res),
                                                               capi.ptr(norms)))
        prm, res = capi.BicgstabParams(1e-8, 0.0, 1000, 0, 0), capi.KrylovResult()
        return res, lambda: capi.check(L.b200_bicgstab_solve(ctx, dA, damg, C.byref(prm), db.ptr, dx.ptr, C.byref(res),
                                                             capi.ptr(norms)))
    for kind in ("pcg", "gmres", "cogmres", "flexgmres", "lgmres", "bicgstab"):
        res, run = solver(kind)

        def step():
            capi.check(L.b200_vec_fill(ctx, n, 0.0, dx.ptr))
            run()
        for _ in range(args.warmup):
            step()
        l0 = L.b200_ctx_launch_count(ctx)
        capi.check(L.b200_timer_start(ctx))
        for _ in range(args.steps):
            step()
        capi.check(L.b200_timer_stop(ctx, C.byref(ms)))
        it = res.num_iterations
        # one BiCGSTAB iteration applies the operator and the preconditioner twice
        print(json.dumps({"solver": kind, "workload": "7-pt Laplacian %d^3, rlx %d" % (args.n, args.rlx), "dof": n,
                          "iterations": it, "ms_per_solve": ms.value / args.steps,
                          "dof_it_per_s": n * it * args.steps / (ms.value * 1e-3),
                          "rel_residual": res.rel_residual_norm, "converged": res.converged,
                          "launches_per_solve": (L.b200_ctx_launch_count(ctx) - l0) / args.steps,
                          "k_dim": args.k if "gmres" in kind else None}))


if __name__ == "__main__":
    main()
