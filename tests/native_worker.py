"""One rank of the native multi-rank mode driven from Python (tests/test_native_host.py): N processes that
did NOT come from mpirun_shim meet through SHIM_DIR (the way torchrun-started bench ranks do), build the
distributed problem with hypre's own generators and run AMG-PCG.  With the device path switched off this is the
reference on N ranks -- the CPU check that hypre_b200/hypre.py drives the MPI build exactly like ij.c does.
argv: P Q R nx ny nz relax_type [27]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["HYPRE_B200_NATIVE"] = "1"
from hypre_b200 import hypre as H  # noqa: E402


def main():
    a = sys.argv[1:]
    grid = tuple(int(v) for v in a[0:3])
    dims = tuple(int(v) for v in a[3:6])
    rlx = int(a[6])
    gen = H.laplacian_27pt if len(a) > 7 and a[7] == "27" else H.laplacian
    rank, size = H.world()
    H.set_device_path(os.environ.get("NATIVE_WORKER_DEVICE", "0") == "1")
    A = gen(*dims, grid=grid)
    b = A.vector(1.0)
    x = A.vector(0.0)
    amg = H.BoomerAMG(relax_type=rlx)
    pcg = H.ParCSRPCG(tol=1e-8, max_iter=100, two_norm=1, print_level=0, logging=1)
    pcg.SetPrecond(amg)
    pcg.Setup(A, b, x)
    pcg.Solve(A, b, x)
    rel = pcg.rel_norms()
    local = float((x.numpy() ** 2).sum())
    if rank == 0:
        print("NATIVE_OK ranks=%d rows=%d levels=%d iterations=%d final=%.6e local_x2=%.9e"
              % (size, A.global_rows, amg.num_levels, pcg.num_iterations, rel[-1] if len(rel) else float("nan"), local), flush=True)
    H.lib().MPI_Finalize()


if __name__ == "__main__":
    main()
