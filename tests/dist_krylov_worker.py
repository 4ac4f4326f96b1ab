"""gloo worker (world_size >= 2, CPU): the GMRES / COGMRES / FlexGMRES / LGMRES / BiCGSTAB loops of
hypre_b200/csrc/krylov_core.hpp on a row-partitioned problem.  Every rank partitions the same global matrix with
b200_rowpart; the operator is the halo exchange of tests/dist_worker.py (torch.distributed point-to-point
messages following the plan) plus the local diag/offd products, inner products are summed with
dist.all_reduce -- the two places where the device backend communicates, too (op_matvec, scalars_allreduce).
Rank 0 compares with the single-rank oracle: same iteration counts, same histories, same solution."""
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from hypre_b200 import capi  # noqa: E402
from dist_worker import exchange, lap7  # noqa: E402
import oracle_util as O  # noqa: E402

MV = C.CFUNCTYPE(None, C.c_double, C.POINTER(C.c_double), C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double))
AR = C.CFUNCTYPE(None, C.POINTER(C.c_double), C.c_int)


def diag_first(A):
    """hypre's storage order for a diag block: the diagonal entry first in its row (the oracle's diagonal scaling
    reads it there, like HYPRE_ParCSRDiagScale)"""
    import scipy.sparse as sp
    A = A.tocsr()
    ind, dat = A.indices.copy(), A.data.copy()
    for r in range(A.shape[0]):
        lo, hi = A.indptr[r], A.indptr[r + 1]
        k = lo + int(np.where(ind[lo:hi] == r)[0][0])
        ind[lo], ind[k] = ind[k], ind[lo]
        dat[lo], dat[k] = dat[k], dat[lo]
    return sp.csr_matrix((dat, ind, A.indptr.copy()), shape=A.shape)


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    L = C.CDLL(os.path.join(ROOT, "tests", "hostbackend", "libkrylov_host_test.so"))
    vp, i = C.c_void_p, C.c_int
    L.kht_gmres_dist.argtypes = [i, vp, MV, AR, vp, vp, vp, vp, vp]
    L.kht_bicgstab_dist.argtypes = [i, vp, MV, AR, vp, vp, vp, vp, vp]
    L.kht_lgmres_dist.argtypes = [i, vp, MV, AR, vp, vp, vp, vp, vp]
    A = lap7(9, 7, 2 * world + 2)
    n = A.shape[0]
    rs = capi.even_starts(n, world)
    part = capi.RowPart(rank, world, A, rs, rs)
    D, Off = part.blocks()
    nl = part.n_rows
    diag = np.ascontiguousarray(A.diagonal()[rs[rank]:rs[rank + 1]])

    def arr(p):
        return np.ctypeslib.as_array(p, shape=(nl,))

    def matvec(alpha, x, beta, b, y):
        xl = arr(x)
        ghost = exchange(part, xl)
        t = D @ xl + (Off @ ghost if part.n_cols_offd else 0.0)
        out = alpha * t
        if beta != 0.0:
            out = out + beta * arr(b)
        arr(y)[:] = out

    def allreduce(vals, cnt):
        v = np.ctypeslib.as_array(vals, shape=(cnt,))
        t = torch.from_numpy(v.copy())
        dist.all_reduce(t)
        v[:] = t.numpy()
    mv_cb, ar_cb = MV(matvec), AR(allreduce)

    bg = 1.0 + 0.01 * np.cos(np.arange(n))
    b = np.ascontiguousarray(bg[rs[rank]:rs[rank + 1]])
    orc = O.OracleAMG.from_arrays([dict(A=diag_first(A), P=None)], [18, 18, 18, 18], [1, 1, 1, 1], 1, 0, 2, 1, None) if rank == 0 else None
    ok = True
    for kind, k_dim, precond in (("gmres", 5, 1), ("gmres", 4, 2), ("cogmres", 6, 1), ("flexgmres", 5, 1), ("lgmres", 6, 1), ("bicgstab", 0, 1),
                                 ("bicgstab", 0, 2)):
        x = np.zeros(nl)
        norms = np.zeros(501)
        res = capi.KrylovResult()
        dptr = capi.ptr(diag) if precond == 1 else None
        if kind == "bicgstab":
            prm = capi.BicgstabParams(1e-8, 0.0, 500, 0, 1 if precond == 1 else 0)
            st = L.kht_bicgstab_dist(nl, dptr, mv_cb, ar_cb, C.byref(prm), capi.ptr(b), capi.ptr(x), C.byref(res), capi.ptr(norms))
        elif kind == "lgmres":
            prm = capi.LgmresParams(1e-8, 0.0, 500, 0, k_dim, 2, 1, 1 if precond == 1 else 0)
            st = L.kht_lgmres_dist(nl, dptr, mv_cb, ar_cb, C.byref(prm), capi.ptr(b), capi.ptr(x), C.byref(res), capi.ptr(norms))
        else:
            prm = capi.GmresParams(1e-8, 0.0, 500, 0, k_dim, 0, 1 if precond == 1 else 0, 1 if kind == "cogmres" else 0,
                                   1 if kind == "flexgmres" else 0)
            st = L.kht_gmres_dist(nl, dptr, mv_cb, ar_cb, C.byref(prm), capi.ptr(b), capi.ptr(x), C.byref(res), capi.ptr(norms))
        xs = [None] * world
        dist.all_gather_object(xs, x)
        if rank == 0:
            kw = dict(precond=precond, max_iter=500)
            if kind != "bicgstab":
                kw["k_dim"] = k_dim
            xo, it, no, rel, conv = getattr(orc, kind)(bg, np.zeros(n), **kw)
            good = (st == 0 and res.num_iterations == it and res.converged == conv == 1 and
                    np.allclose(norms[:it + 1], no, rtol=1e-7, atol=1e-13 * no[0]) and
                    np.allclose(np.concatenate(xs), xo, rtol=1e-7, atol=1e-10))
            if not good:
                print("MISMATCH", kind, k_dim, precond, st, res.num_iterations, it, res.converged, conv,
                      np.max(np.abs(norms[:it + 1] - no[:it + 1]) / no[:it + 1]) if res.num_iterations == it else None)
            ok &= bool(good)
    part.destroy()
    flag = torch.tensor([1.0 if ok else 0.0], dtype=torch.float64)
    dist.broadcast(flag, 0)
    if rank == 0:
        print("DIST_KRYLOV_OK" if ok else "DIST_KRYLOV_FAIL")
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1.0 else 1)


if __name__ == "__main__":
    main()
