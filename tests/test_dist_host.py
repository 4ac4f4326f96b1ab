"""CPU tests of the N>1 host logic: b200_rowpart (row partition + halo plan, the counterpart of
GenerateDiagAndOffd / hypre_MatvecCommPkgCreate) in a single-process emulation over several rank
counts, a real 2-process gloo run of tests/dist_worker.py, and a comparison with what N real ranks of
the reference (oracle/_ref/dump_parcsr over hypre_b200/host/mpishim) build and negotiate for a whole hierarchy."""
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


# ------------------------------------------------------------------ against the reference on N real ranks
REF_BIN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref")
HAVE_MP = all(os.path.exists(os.path.join(REF_BIN, f)) for f in ("mpirun_shim", "dump_parcsr"))


def read_dump(path):
    """one rank's file written by oracle/dump_parcsr.c"""
    raw = np.fromfile(path, dtype=np.int32)
    h = raw[:12]
    (grows, gcols, first_row, n, first_col, ncd, nco, nnzd, nnzo, ns, nse, nr) = [int(v) for v in h]
    pos = [12]

    def take(k):
        a = raw[pos[0]:pos[0] + k].copy(); pos[0] += k
        return a
    d = dict(global_rows=grows, global_cols=gcols, first_row=first_row, n_rows=n, first_col=first_col, n_cols_diag=ncd)
    d["diag_i"], d["diag_j"] = take(n + 1), take(nnzd)
    d["offd_i"], d["offd_j"] = take(n + 1), take(nnzo)
    d["col_map_offd"] = take(nco)
    d["send_procs"], d["send_map_starts"], d["send_map_elmts"] = take(ns), take(ns + 1), take(nse)
    d["recv_procs"], d["recv_vec_starts"] = take(nr), take(nr + 1)
    vals = raw[pos[0]:].view(np.float64)
    d["diag_a"], d["offd_a"] = vals[:nnzd].copy(), vals[nnzd:nnzd + nnzo].copy()
    assert len(vals) == nnzd + nnzo
    return d


def global_from_dumps(parts):
    """global CSR in the reference's numbering; every row = its diag entries, then its offd entries, each in
    the reference's storage order"""
    ii, jj, aa = [0], [], []
    for d in parts:
        for r in range(d["n_rows"]):
            dk = slice(d["diag_i"][r], d["diag_i"][r + 1]); ok = slice(d["offd_i"][r], d["offd_i"][r + 1])
            jj += list(d["diag_j"][dk] + d["first_col"]) + list(d["col_map_offd"][d["offd_j"][ok]])
            aa += list(d["diag_a"][dk]) + list(d["offd_a"][ok])
            ii.append(len(jj))
    shape = (parts[0]["global_rows"], parts[0]["global_cols"])
    return sp.csr_matrix((np.array(aa), np.array(jj, dtype=np.int32), np.array(ii, dtype=np.int32)), shape=shape)


@pytest.mark.skipif(not HAVE_MP, reason="oracle/_ref multi-rank build absent (make -C oracle)")
@pytest.mark.parametrize("pts,n,P", [(7, (12, 10, 8), (2, 2, 1)), (27, (9, 8, 7), (1, 3, 1)), (7, (6, 6, 9), (1, 1, 2)),
                                     (7, (16, 12, 10), (2, 2, 2))])
def test_rowpart_equals_the_reference_on_real_ranks(tmp_path, pts, n, P):
    """b200_rowpart_create (no communication) against what N ranks of the reference build and negotiate for
    every operator of a BoomerAMG hierarchy -- A_l and the rectangular P_l: diag/offd blocks, col_map_offd
    and the whole hypre_ParCSRCommPkg (send_procs, send_map_starts, send_map_elmts, recv_procs,
    recv_vec_starts) must be identical, rank by rank."""
    nranks = P[0] * P[1] * P[2]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    p = subprocess.run([os.path.join(REF_BIN, "mpirun_shim"), "-np", str(nranks), os.path.join(REF_BIN, "dump_parcsr"),
                        str(pts)] + [str(v) for v in n + P] + [str(tmp_path)], capture_output=True, text=True, timeout=300, env=env)
    assert p.returncode == 0 and "DUMP_OK" in p.stdout, p.stderr[-500:]
    levels = int(p.stdout.split("levels")[1].split()[0])
    checked = 0
    for name in ["A%d" % l for l in range(levels)] + ["P%d" % l for l in range(levels - 1)]:
        parts = [read_dump(os.path.join(str(tmp_path), "%s.%d.bin" % (name, r))) for r in range(nranks)]
        A = global_from_dumps(parts)
        rs = [d["first_row"] for d in parts] + [parts[0]["global_rows"]]
        cs = [d["first_col"] for d in parts] + [parts[0]["global_cols"]]
        for r, d in enumerate(parts):
            rp = capi.RowPart(r, nranks, A, rs, cs)
            for key in ("diag_i", "diag_j", "offd_i", "offd_j", "col_map_offd", "recv_procs", "send_procs",
                        "send_map_elmts"):
                assert np.array_equal(getattr(rp, key), d[key]), (name, r, key)
            # (the reference leaves a one-element starts array when a rank has no neighbours)
            assert np.array_equal(rp.send_map_starts, d["send_map_starts"]), (name, r)
            if len(d["recv_procs"]):
                assert np.array_equal(rp.recv_vec_starts, d["recv_vec_starts"]), (name, r)
            assert np.array_equal(rp.diag_a, d["diag_a"]) and np.array_equal(rp.offd_a, d["offd_a"]), (name, r)
            rp.destroy()
            checked += 1
    assert checked == nranks * (2 * levels - 1)
