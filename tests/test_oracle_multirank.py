"""The plain-C oracle against the reference running on N real ranks.

oracle/_ref/dump_parcsr (our driver over the reference's MPI build + hypre_b200/host/mpishim) does what
`mpirun -np N ij -solver 1 -rlx <type> [-rhsrand] -n .. -P ..` does and writes the distributed hierarchy,
the smoother data, the right-hand side and PCG's logged norms.  The test gathers the hierarchy into global
matrices (the reference's numbering), hands them to the serial oracle restatement (oracle/amg_oracle*.c;
rank-local strict lower triangles for two-stage GS) and requires the same iteration count and the same
per-iteration residual norms to 1e-9 -- for configurations whose final residual the reference's own
regression suite records (src/test/TEST_ij/smoother.saved), and for the two-stage Gauss-Seidel smoothers,
which have no golden there."""
import os
import subprocess

import numpy as np
import pytest
import scipy.sparse as sp

import oracle_util as O
from test_dist_host import HAVE_MP, REF_BIN, global_from_dumps, read_dump

pytestmark = pytest.mark.skipif(not HAVE_MP, reason="oracle/_ref multi-rank build absent (make -C oracle)")

NORM_TOL = 1e-9

# (points, n, P, relax, rhsrand, golden (iterations, final) from TEST_ij/smoother.saved or None,
#  extra driver arguments: cycle_type (2 = W), num_sweeps, aggressive-coarsening levels)
CONFIGS = [
    (7, (20, 20, 10), (2, 2, 1), 16, 1, (6, 2.510130e-09), ()),     # smoother.jobs:72  -> smoother.out.12
    (7, (40, 40, 20), (2, 2, 1), 16, 0, (7, 7.310897e-09), ()),     # smoother.jobs:102 -> smoother.out.22
    (7, (20, 20, 10), (2, 2, 1), 18, 1, None, ()),
    (7, (15, 30, 10), (1, 3, 1), 18, 1, None, ()),
    (27, (12, 12, 12), (2, 1, 2), 11, 0, None, ()),
    (27, (12, 12, 12), (2, 2, 2), 12, 1, None, ()),
    (7, (24, 24, 24), (2, 2, 2), 7, 0, None, ()),
    (7, (20, 20, 20), (2, 2, 1), 18, 1, None, (2, 1, 0)),           # W-cycle
    (27, (14, 14, 14), (1, 2, 2), 16, 0, None, (1, 2, 0)),          # two sweeps
    (7, (24, 24, 12), (2, 2, 1), 18, 1, None, (1, 1, 1)),           # one level of aggressive coarsening
    (7, (20, 20, 20), (3, 1, 1), 11, 1, None, (2, 2, 0)),           # W-cycle, two sweeps, two-stage GS, 3 ranks
]


def f64(path):
    return np.fromfile(path, dtype=np.float64)


@pytest.mark.parametrize("pts,n,P,rlx,rhsrand,golden,extra", CONFIGS)
def test_oracle_repeats_the_multirank_reference_solve(tmp_path, pts, n, P, rlx, rhsrand, golden, extra):
    nranks = P[0] * P[1] * P[2]
    d = str(tmp_path)
    env = dict(os.environ, OMP_NUM_THREADS="1")
    p = subprocess.run([os.path.join(REF_BIN, "mpirun_shim"), "-np", str(nranks), os.path.join(REF_BIN, "dump_parcsr"),
                        str(pts)] + [str(v) for v in n + P] + [d, str(rlx), str(rhsrand)] + [str(v) for v in extra],
                       capture_output=True, text=True, timeout=600, env=env)
    assert p.returncode == 0 and "DUMP_OK" in p.stdout, p.stderr[-500:]
    meta = {ln.split()[0]: ln.split()[1:] for ln in open(os.path.join(d, "meta.txt"))}
    nl, iters = int(meta["levels"][0]), int(meta["iterations"][0])
    ref_rel = f64(os.path.join(d, "rel_norms.f64"))[1:iters + 1]
    if golden:
        assert iters == golden[0] and abs(float(meta["final_rel"][0]) - golden[1]) <= 2e-6 * golden[1]

    levels, starts = [], []
    for l in range(nl):
        partsA = [read_dump(os.path.join(d, "A%d.%d.bin" % (l, r))) for r in range(nranks)]
        lv = {"A": global_from_dumps(partsA), "P": None, "weight": float(meta["relax_weight"][l])}
        if l < nl - 1:
            lv["P"] = global_from_dumps([read_dump(os.path.join(d, "P%d.%d.bin" % (l, r))) for r in range(nranks)])
        lv["l1"] = np.concatenate([f64(os.path.join(d, "l1_%d.%d.f64" % (l, r))) for r in range(nranks)])
        lv["ds"] = np.concatenate([f64(os.path.join(d, "ds_%d.%d.f64" % (l, r))) for r in range(nranks)])
        lv["coefs"] = f64(os.path.join(d, "coefs_%d.f64" % l))
        levels.append(lv)
        starts.append([q["first_row"] for q in partsA] + [partsA[0]["global_rows"]])
    dense = None
    if os.path.exists(os.path.join(d, "dense.f64")):
        dense = f64(os.path.join(d, "dense.f64"))
        N = levels[-1]["A"].shape[0]
        assert dense.size == N * N
        dense = dense.reshape(N, N)
    orc = O.OracleAMG.from_arrays(levels, [int(v) for v in meta["relax_type"]], [int(v) for v in meta["num_sweeps"]],
                                  int(meta["cycle_type"][0]), int(meta["fcycle"][0]), int(meta["cheby_order"][0]),
                                  int(meta["cheby_scale"][0]), dense)
    if rlx in (11, 12):
        for l in range(nl):
            orc.set_level_blocks(l, starts[l])
    b = np.concatenate([f64(os.path.join(d, "b.%d.f64" % r)) for r in range(nranks)])
    xref = np.concatenate([f64(os.path.join(d, "x.%d.f64" % r)) for r in range(nranks)])
    x, it, rel, conv = orc.pcg(b, np.zeros_like(b), tol=1e-8, max_iter=1000)
    assert it == iters, (it, iters)
    np.testing.assert_allclose(rel, ref_rel, rtol=NORM_TOL)
    np.testing.assert_allclose(x, xref, rtol=1e-8, atol=1e-11 * np.max(np.abs(xref)))
    orc.destroy()
