#!/usr/bin/env python
"""Level-0 operator timings on N ranks (torchrun) or 1 rank: SpMV / l1-Jacobi / Chebyshev with the
halo exchange inside, CUDA events over `reps` back-to-back launches, max over ranks."""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gen", default="laplacian")
    ap.add_argument("--dims", type=int, nargs=3, default=[256, 256, 256])
    ap.add_argument("--reps", type=int, default=50)
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from hypre_b200 import capi, hypre as H
    L, HL = capi.lib(), H.lib()
    if world > 1:
        uid = C.create_string_buffer(128)
        if rank == 0:
            capi.check(L.b200_nccl_unique_id(uid))
        t = torch.tensor(list(uid.raw), dtype=torch.uint8, device="cuda")
        dist.broadcast(t, 0)
        uid = C.create_string_buffer(bytes(t.cpu().tolist()), 128)
        assert HL.hypre_b200_init_dist(local, rank, world, uid) == 0
    ctx = C.c_void_p(HL.hypre_b200_context())
    A = getattr(H, a.gen)(*a.dims)
    dA = A.device_mirror()
    n = C.c_int(); nnz = C.c_longlong()
    capi.check(L.b200_parcsr_info(dA, C.byref(n), None, C.byref(nnz), None, None))
    n, nnz = n.value, nnz.value
    v = [capi.DeviceVector(ctx, n, np.sin(np.arange(n)) * 1e-3 + 1.0) for _ in range(7)]
    coefs = np.array([0.9, -0.3, 0.05])
    ms = C.c_float()
    spmv = 12.0 * nnz + 4.0 * n

    def timed(fn, bytes_):
        for _ in range(5):
            fn()
        capi.check(L.b200_ctx_sync(ctx))
        if world > 1:
            dist.barrier(); torch.cuda.synchronize()
        capi.check(L.b200_timer_start(ctx))
        for _ in range(a.reps):
            fn()
        capi.check(L.b200_timer_stop(ctx, C.byref(ms)))
        t = ms.value / a.reps
        if world > 1:
            tt = torch.tensor([t], device="cuda"); dist.all_reduce(tt, op=dist.ReduceOp.MAX); t = tt.item()
        return {"us": round(t * 1e3, 1), "GB/s_per_gpu": round(bytes_ / t / 1e6, 0)}

    out = {"ranks": world, "dims": a.dims, "rows_per_rank": n, "nnz_per_rank": nnz, "halo": os.environ.get("B200_HALO", "p2p")}
    out["spmv"] = timed(lambda: capi.check(L.b200_parcsr_matvec(dA, 1.0, v[0].ptr, 0.0, None, v[1].ptr)), spmv + 16.0 * n)
    out["cheby2"] = timed(lambda: capi.check(L.b200_relax_cheby(dA, v[2].ptr, v[3].ptr, capi.ptr(coefs), 2, 1, v[0].ptr,
                                                               v[4].ptr, v[5].ptr, v[6].ptr, 0)), 2 * (spmv + 40.0 * n))
    out["axpy"] = timed(lambda: capi.check(L.b200_vec_axpy(ctx, n, 0.5, v[0].ptr, v[1].ptr)), 24.0 * n)
    r = C.c_double()
    out["dot_allreduce_sync"] = timed(lambda: capi.check(L.b200_vec_dot(ctx, n, v[0].ptr, v[1].ptr, C.byref(r))), 16.0 * n)
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
