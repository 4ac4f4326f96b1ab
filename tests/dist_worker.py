"""gloo worker (world_size >= 2, CPU): every rank partitions the same global matrix with
b200_rowpart, performs the halo exchange of hypre_ParCSRMatrixMatvec (job 1) and MatvecT
(job 2) with torch.distributed point-to-point messages following the plan, and rank 0 checks the
assembled result against the global product.  Exercises the N>1 host logic without a GPU."""
import os
import sys

import numpy as np
import scipy.sparse as sp
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hypre_b200 import capi  # noqa: E402


def lap7(nx, ny, nz):
    def t(n):
        return sp.diags([-np.ones(n - 1), 2 * np.ones(n), -np.ones(n - 1)], [-1, 0, 1])
    ex, ey, ez = sp.identity(nx), sp.identity(ny), sp.identity(nz)
    return (sp.kron(ez, sp.kron(ey, t(nx))) + sp.kron(ez, sp.kron(t(ny), ex)) + sp.kron(t(nz), sp.kron(ey, ex))).tocsr()


def exchange(part, x_local, reverse=False, ghost_out=None):
    """job 1: returns the ghost vector.  job 2 (reverse): sends ghost_out to the owners and returns
    the received contributions in send_map order."""
    reqs, keep = [], []
    if not reverse:
        ghost = torch.zeros(part.n_cols_offd, dtype=torch.float64)
        for i, q in enumerate(part.recv_procs):
            buf = ghost[part.recv_vec_starts[i]:part.recv_vec_starts[i + 1]]
            reqs.append(dist.irecv(buf, src=int(q)))
        packed = torch.from_numpy(x_local[part.send_map_elmts].copy())
        for i, q in enumerate(part.send_procs):
            buf = packed[part.send_map_starts[i]:part.send_map_starts[i + 1]].clone()
            keep.append(buf)
            reqs.append(dist.isend(buf, dst=int(q)))
        for r in reqs:
            r.wait()
        return ghost.numpy()
    recv = torch.zeros(part.n_send_elmts, dtype=torch.float64)
    for i, q in enumerate(part.send_procs):
        buf = recv[part.send_map_starts[i]:part.send_map_starts[i + 1]]
        reqs.append(dist.irecv(buf, src=int(q)))
    go = torch.from_numpy(np.ascontiguousarray(ghost_out))
    for i, q in enumerate(part.recv_procs):
        buf = go[part.recv_vec_starts[i]:part.recv_vec_starts[i + 1]].clone()
        keep.append(buf)
        reqs.append(dist.isend(buf, dst=int(q)))
    for r in reqs:
        r.wait()
    return recv.numpy()


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    A = lap7(9, 7, 2 * world + 1)
    n = A.shape[0]
    # a rectangular "interpolation" operator: coarse point = every second fine point
    cpts = np.arange(0, n, 2)
    nc = len(cpts)
    rows = np.concatenate([cpts, np.arange(1, n, 2), np.arange(1, n - 1, 2)])
    cols = np.concatenate([np.arange(nc), np.arange(1, n, 2) // 2, np.arange(1, n - 1, 2) // 2 + 1])
    vals = np.concatenate([np.ones(nc), 0.5 * np.ones(len(np.arange(1, n, 2))), 0.5 * np.ones(len(np.arange(1, n - 1, 2)))])
    P = sp.csr_matrix((vals, (rows, cols)), shape=(n, nc))
    rs = capi.even_starts(n, world)
    cs_c = np.array([np.searchsorted(cpts, s) for s in rs], dtype=np.int32)   # coarse points inherit the owner
    rng = np.random.default_rng(5)
    x = rng.standard_normal(n)
    xc = rng.standard_normal(nc)
    ok = True
    for M, r_st, c_st, xin, xT in ((A, rs, rs, x, x), (P, rs, cs_c, xc, x)):
        part = capi.RowPart(rank, world, M, r_st, c_st)
        D, O = part.blocks()
        # ---- Matvec: y = M x
        x_loc = xin[c_st[rank]:c_st[rank + 1]]
        ghost = exchange(part, x_loc)
        assert np.array_equal(ghost, xin[part.col_map_offd])
        y_loc = D @ x_loc + (O @ ghost if part.n_cols_offd else 0.0)
        # ---- MatvecT: z = M^T x  (job 2 exchange + add at send_map_elmts)
        xr = xT[r_st[rank]:r_st[rank + 1]]
        z_loc = D.T @ xr
        contrib = exchange(part, None, reverse=True, ghost_out=(O.T @ xr) if part.n_cols_offd else np.zeros(0))
        np.add.at(z_loc, part.send_map_elmts, contrib)
        ys = [None] * world
        zs = [None] * world
        dist.all_gather_object(ys, y_loc)
        dist.all_gather_object(zs, z_loc)
        if rank == 0:
            y = np.concatenate(ys); z = np.concatenate(zs)
            ok &= np.allclose(y, M @ xin, rtol=1e-13, atol=1e-13)
            ok &= np.allclose(z, M.T @ xT, rtol=1e-13, atol=1e-13)
        part.destroy()
    # allreduce of the PCG dot products
    t = torch.tensor([float(np.dot(x[rs[rank]:rs[rank + 1]], x[rs[rank]:rs[rank + 1]]))], dtype=torch.float64)
    dist.all_reduce(t)
    ok &= abs(t.item() - float(np.dot(x, x))) < 1e-10
    if rank == 0:
        print("DIST_OK" if ok else "DIST_FAIL")
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
