"""N-GPU parity worker (torchrun, one process per GPU): the distributed device PCG (row-partitioned
hierarchy, NCCL halo exchange + allreduce) against the single-rank plain-C oracle on the same
hierarchy.  l1-Jacobi and Chebyshev do not depend on the partition; two-stage GS uses the strict lower part
of the rank-local diag block (par_relax.c:1505-1588), so the oracle is told the row partition of
every distributed level.  In all cases the per-iteration residual norms must agree to 1e-9 and the
iteration counts exactly."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    os.environ.setdefault("OMP_NUM_THREADS", str(max(1, (os.cpu_count() or 1) // world)))
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from hypre_b200 import capi, hypre as H
    import oracle_util as O
    L, HL = capi.lib(), H.lib()
    uid = C.create_string_buffer(128)
    if rank == 0:
        capi.check(L.b200_nccl_unique_id(uid))
    t = torch.tensor(list(uid.raw), dtype=torch.uint8, device="cuda")
    dist.broadcast(t, 0)
    uid = C.create_string_buffer(bytes(t.cpu().tolist()), 128)
    assert HL.hypre_b200_init_dist(local, rank, world, uid) == 0
    ok = True
    for name in ["lap7_32_rlx18", "lap7_24_rlx16", "lap27_20_rlx11", "lap27_20_rlx12", "lap7_64_rlx16",
                 "rot_512_rlx18_agg1"]:
        A, b, x, amg, pcg = O.build_case(H, name)
        n = A.global_rows
        H.set_device_path(True)
        x.numpy()[:] = 0.0
        H.clear_errors()
        pcg.Solve(A, b, x)                      # each rank solves its row block; x holds this rank's slice
        assert H.error_flag() == 0, H.error_flag()
        it, rel = pcg.num_iterations, pcg.rel_norms()
        first = HL.hypre_b200_first_row(amg.h)
        dA = C.c_void_p(); nloc = C.c_int()
        capi.check(L.b200_amg_level_matrix(amg.device_mirror(), 0, C.byref(dA)))
        capi.check(L.b200_parcsr_info(dA, C.byref(nloc), None, None, None, None))
        xs = torch.zeros(n, dtype=torch.float64, device="cuda")
        xs[first:first + nloc.value] = torch.from_numpy(x.numpy()[first:first + nloc.value].copy()).cuda()
        dist.all_reduce(xs)
        xg = xs.cpu().numpy()
        orc = O.OracleAMG(amg)
        # give the oracle the row partition of every distributed level: two-stage GS uses the strict
        # lower triangle of each rank's diag block (par_relax.c:1562-1581), everything else is
        # partition-independent
        st = np.zeros(world + 1, dtype=np.int32)
        for l in range(amg.num_levels):
            if HL.hypre_b200_amg_level_starts(amg.h, l, capi.ptr(st)) == world:
                orc.set_level_blocks(l, st)
        xo, ito, relo, conv = orc.pcg(b.numpy(), np.zeros(n))
        orc.destroy()
        good = it == ito and np.allclose(rel, relo, rtol=1e-9, atol=0) and \
            np.max(np.abs(xg - xo)) <= 1e-8 * np.max(np.abs(xo))
        # the assembled solution really solves the global system
        Ad = A.blocks()[0]
        res = np.linalg.norm(b.numpy() - Ad @ xg) / np.linalg.norm(b.numpy())
        good = good and res < 1.05e-8
        if rank == 0:
            print("%-20s ranks %d  iterations %d (oracle %d)  final %.6e (oracle %.6e)  true residual %.3e  %s"
                  % (name, world, it, ito, rel[-1], relo[-1], res, "OK" if good else "MISMATCH"), flush=True)
        ok = ok and good
    flag = torch.tensor([0 if ok else 1], device="cuda")
    dist.all_reduce(flag)
    if rank == 0:
        print("DIST_GPU_OK" if flag.item() == 0 else "DIST_GPU_FAIL", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 0 else 1)


if __name__ == "__main__":
    main()
