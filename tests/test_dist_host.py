"""CPU tests of the N>1 host logic: b200_rowpart (row partition + halo plan, the counterpart of
GenerateDiagAndOffd / hypre_MatvecCommPkgCreate) in a single-process emulation over several rank
counts, and a real 2-process gloo run of tests/dist_worker.py."""
import os
import subprocess
import sys

import numpy as np
import pytest
import scipy.sparse as sp

from hypre_b200 import ROOT, capi


def emulate_matvec(A, nranks, rs, cs, x):
    """every rank's blocks + plan; halo exchange emulated by indexing through the plan"""
    parts = [capi.RowPart(r, nranks, A, rs, cs) for r in range(nranks)]
    y = np.zeros(A.shape[0])
    z = np.zeros(A.shape[1])
    xr = np.arange(A.shape[0]) * 0.25 + 1.0
    sendbuf = {}
    for r, p in enumerate(parts):
        x_loc = x[cs[r]:cs[r + 1]]
        for i, q in enumerate(p.send_procs):
            sendbuf[(r, int(q))] = x_loc[p.send_map_elmts[p.send_map_starts[i]:p.send_map_starts[i + 1]]]
    for r, p in enumerate(parts):
        D, O = p.blocks()
        ghost = np.zeros(p.n_cols_offd)
        for i, q in enumerate(p.recv_procs):
            msg = sendbuf[(int(q), r)]
            assert len(msg) == p.recv_vec_starts[i + 1] - p.recv_vec_starts[i]
            ghost[p.recv_vec_starts[i]:p.recv_vec_starts[i + 1]] = msg
        assert np.array_equal(ghost, x[p.col_map_offd])
        assert np.all(np.diff(p.col_map_offd) > 0)
        y[rs[r]:rs[r + 1]] = D @ x[cs[r]:cs[r + 1]] + (O @ ghost if p.n_cols_offd else 0.0)
        # transpose product: local part + contributions routed through the plan
        z[cs[r]:cs[r + 1]] += D.T @ xr[rs[r]:rs[r + 1]]
        if p.n_cols_offd:
            z[p.col_map_offd] += O.T @ xr[rs[r]:rs[r + 1]]
    for p in parts:
        p.destroy()
    return y, z, xr


@pytest.mark.parametrize("nranks", [1, 2, 3, 5, 8])
def test_rowpart_matvec_equals_global(H, nranks):
    H.set_device_path(False)
    A = H.laplacian_27pt(7, 6, 9)
    Ad = A.blocks()[0]
    n = Ad.shape[0]
    rs = capi.even_starts(n, nranks)
    x = np.sin(np.arange(n)) + 2.0
    y, z, xr = emulate_matvec(Ad, nranks, rs, rs, x)
    np.testing.assert_allclose(y, Ad @ x, rtol=1e-13)
    np.testing.assert_allclose(z, Ad.T @ xr, rtol=1e-13)
    A.destroy()
    H.set_device_path(True)


def test_rowpart_keeps_diagonal_first_and_handles_empty_ranks(H):
    """diag rows keep hypre's storage order (diagonal entry first); ranks may own zero rows"""
    H.set_device_path(False)
    A = H.laplacian(5, 4, 3)
    Ad = A.blocks()[0]
    n = Ad.shape[0]
    rs = np.array([0, 0, 25, 25, n], dtype=np.int32)      # ranks 0 and 2 are empty
    for r in range(4):
        p = capi.RowPart(r, 4, Ad, rs, rs)
        assert p.n_rows == rs[r + 1] - rs[r]
        for i in range(p.n_rows):
            assert p.diag_j[p.diag_i[i]] == i              # first stored entry of a row is its diagonal
        if p.n_rows == 0:
            assert p.num_sends == 0 and p.num_recvs == 0 and p.n_cols_offd == 0
        p.destroy()
    x = np.cos(np.arange(n))
    y, z, xr = emulate_matvec(Ad, 4, rs, rs, x)
    np.testing.assert_allclose(y, Ad @ x, rtol=1e-13)
    A.destroy()
    H.set_device_path(True)


def test_rowpart_rectangular_interpolation(H):
    """P of a real hierarchy with the inherited coarse partition (C points stay with their owner)"""
    import oracle_util as O
    H.set_device_path(False)
    A, b, x, amg, pcg = O.build_case(H, "lap7_24_rlx16")
    P = amg.P(0).blocks()[0]
    nf, nc = P.shape
    nranks = 3
    rs = capi.even_starts(nf, nranks)
    # C points = rows of P with the single entry 1.0 (identity rows)
    is_c = (np.diff(P.indptr) == 1) & (P.data[P.indptr[:-1]] == 1.0)
    assert is_c.sum() == nc
    cs = np.array([is_c[:s].sum() for s in rs], dtype=np.int32)
    xc = np.sin(np.arange(nc)) + 1.5
    y, z, xr = emulate_matvec(P, nranks, rs, cs, xc)
    np.testing.assert_allclose(y, P @ xc, rtol=1e-13)
    np.testing.assert_allclose(z, P.T @ xr, rtol=1e-13)
    H.set_device_path(True)


def test_rowpart_rejects_bad_partitions():
    L = capi.lib()
    A = sp.identity(4, format="csr")
    with pytest.raises(capi.B200Error):
        capi.RowPart(0, 2, A, [0, 2, 3], [0, 2, 4])     # rows do not cover the matrix
    with pytest.raises(capi.B200Error):
        capi.RowPart(2, 2, A, [0, 2, 4], [0, 2, 4])     # rank out of range
    assert b"bad argument" in L.b200_last_error()


def test_two_process_gloo_halo_exchange(built):
    """world_size 2 over gloo: the plan drives real point-to-point messages (tests/dist_worker.py)"""
    env = dict(os.environ, OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.join(ROOT, "tests", "dist_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0 and "DIST_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]
