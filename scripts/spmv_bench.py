#!/usr/bin/env python
"""BASELINE.json configs[1]: stand-alone ParCSR SpMV microbench, 27-pt Laplacian 200^3 (8M rows, ~216M nnz),
hypre_ParCSRMatrixMatvec (src/test/ij.c:4262-4296: y = A x repeated) on 1 / 2 / 4 / 8 B200.

    python scripts/spmv_bench.py                                          # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/spmv_bench.py

Every rank mirrors its 1-D row block of the (replicated, single-rank) host matrix; the halo exchange is inside
the kernel (peer-memory stores over NVLink).  Timing: CUDA events over `reps` back-to-back launches on the
library's stream, max over ranks.  Parity: this rank's rows of y against the CSR row sums of the host matrix
(scipy, the reference's loop order is irrelevant at 1e-12 relative per row scaled by sum |a_ij x_j|).
One JSON line on rank 0."""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gen", default="laplacian_27pt")
    ap.add_argument("--dims", type=int, nargs=3, default=[200, 200, 200])
    ap.add_argument("--reps", type=int, default=50)
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    os.environ["OMP_NUM_THREADS"] = str(max(1, (os.cpu_count() or 1) // world))
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
    vfmt = C.c_int(); st_b = C.c_longlong(); csr_b = C.c_longlong(); ndict = C.c_int()
    capi.check(L.b200_parcsr_storage(dA, C.byref(vfmt), C.byref(st_b), C.byref(csr_b), C.byref(ndict)))
    ng = A.global_rows
    first = HL.hypre_b200_first_row(A.h) if world > 1 else 0
    xg = np.sin(np.arange(ng) * 0.001) + 1.5
    dx = capi.DeviceVector(ctx, n, xg[first:first + n])
    dy = capi.DeviceVector(ctx, n)
    ms = C.c_float()

    def spmv():
        capi.check(L.b200_parcsr_matvec(dA, 1.0, dx.ptr, 0.0, None, dy.ptr))
    spmv()
    # parity on this rank's rows
    Ad, _ = A.blocks()
    rows = Ad[first:first + n]
    ref = rows.dot(xg)
    scale = abs(rows).dot(np.abs(xg)) + 1e-300
    err = float(np.max(np.abs(dy.get() - ref) / scale))
    for _ in range(5):
        spmv()
    capi.check(L.b200_ctx_sync(ctx))
    if world > 1:
        dist.barrier(); torch.cuda.synchronize()
    capi.check(L.b200_timer_start(ctx))
    for _ in range(a.reps):
        spmv()
    capi.check(L.b200_timer_stop(ctx, C.byref(ms)))
    t = ms.value * 1e-3 / a.reps

    def red(v, op):
        if world == 1:
            return v
        tt = torch.tensor([float(v)], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=op)
        return float(tt.item())
    if world > 1:
        t = red(t, dist.ReduceOp.MAX); err = red(err, dist.ReduceOp.MAX)
        nnz_all = red(nnz, dist.ReduceOp.SUM); stored_all = red(st_b.value, dist.ReduceOp.SUM)
    else:
        nnz_all, stored_all = nnz, st_b.value
    csr_bytes = 12.0 * nnz_all + 4.0 * ng + 16.0 * ng
    if rank == 0:
        peak = 6650.0
        try:
            peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        except Exception:
            pass
        print(json.dumps({
            "metric": "ParCSR SpMV rows/s (hypre_ParCSRMatrixMatvec)", "workload": "%s %s, y = A x" % (a.gen, "x".join(map(str, a.dims))),
            "n_gpus": world, "rows": ng, "nnz": int(nnz_all), "us_per_spmv": t * 1e6, "value": ng / t, "unit": "rows/s",
            "csr_gbs": csr_bytes / t / 1e9, "csr_frac_of_peak": csr_bytes / t / 1e9 / (peak * world),
            "stored_gbs": (stored_all + 16.0 * ng) / t / 1e9, "stored_frac_of_peak": (stored_all + 16.0 * ng) / t / 1e9 / (peak * world),
            "value_format": vfmt.value, "n_dictionary": ndict.value, "max_row_err": err, "row_tolerance": 1e-12,
            "parity_ok": bool(err <= 1e-12), "peak_gbs_per_gpu": peak, "reps": a.reps,
            "halo": os.environ.get("B200_HALO", "peer-memory stores fused into the kernel") if world > 1 else None}), flush=True)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
