"""Test-only helpers: ctypes binding of oracle/_ref/liboracle.so (the plain-C restatement)
and problem builders shared by the CPU and GPU tests."""
import ctypes as C
import json
import os

import numpy as np

from hypre_b200 import LIB_ORACLE, ROOT, capi

GOLDEN = os.path.join(ROOT, "tests", "golden", "pcg_histories.json")
_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_ORACLE)
        vp, i, d = C.c_void_p, C.c_int, C.c_double
        L.orc_csr_matvec.argtypes = [i, vp, vp, vp, d, vp, d, vp, vp]
        L.orc_csr_matvecT.argtypes = [i, i, vp, vp, vp, d, vp, d, vp]
        L.orc_relax_l1jacobi.argtypes = [i, vp, vp, vp, vp, vp, d, vp, vp, i]
        L.orc_relax_cheby.argtypes = [i, vp, vp, vp, vp, vp, vp, i, i, vp, vp, vp, vp, vp]
        L.orc_relax_gs2.argtypes = [i, vp, vp, vp, vp, d, i, vp, vp]
        L.orc_gselim.argtypes = [vp, vp, i]
        L.orc_amg_create.restype = vp
        L.orc_amg_create.argtypes = [i]
        L.orc_amg_destroy.argtypes = [vp]
        L.orc_amg_set_level.argtypes = [vp, i, i, vp, vp, vp, i, vp, vp, vp, vp, vp, vp, d]
        L.orc_amg_set_cycle.argtypes = [vp, vp, vp, i, i, i, i]
        L.orc_amg_set_coarse_dense.argtypes = [vp, i, vp]
        L.orc_amg_set_level_blocks.argtypes = [vp, i, i, vp]
        L.orc_relax_gs2_blocks.argtypes = [i, vp, vp, vp, vp, d, i, vp, vp, i, vp]
        L.orc_amg_cycle.argtypes = [vp, vp, vp, i]
        L.orc_pcg.argtypes = [vp, vp, vp, d, i, i, vp, vp, vp]
        L.orc_amg_set_precond_cycles.argtypes = [vp, i]
        L.orc_amg_set_level_cf.argtypes = [vp, i, vp]
        L.orc_amg_set_relax_order.argtypes = [vp, i]
        L.orc_gmres.argtypes = [vp, i, vp, vp, d, d, i, i, i, i, vp, vp, vp]
        L.orc_cogmres.argtypes = [vp, i, vp, vp, d, d, i, i, i, i, vp, vp, vp]
        L.orc_flexgmres.argtypes = [vp, i, vp, vp, d, d, i, i, i, i, vp, vp, vp]
        L.orc_lgmres.argtypes = [vp, i, vp, vp, d, d, i, i, i, i, i, vp, vp, vp]
        L.orc_krylov_set_options.argtypes = [d, i]
        L.orc_bicgstab.argtypes = [vp, i, vp, vp, d, d, i, i, vp, vp, vp]
        _lib = L
    return _lib


p = capi.ptr


def csr_arrays(A):
    A = A.tocsr()
    return (np.ascontiguousarray(A.indptr, dtype=np.int32), np.ascontiguousarray(A.indices, dtype=np.int32),
            np.ascontiguousarray(A.data, dtype=np.float64))


def matvec(A, alpha, x, beta, b):
    ii, jj, aa = csr_arrays(A)
    y = np.zeros(A.shape[0])
    bb = np.ascontiguousarray(b, dtype=np.float64) if b is not None else None
    lib().orc_csr_matvec(A.shape[0], p(ii), p(jj), p(aa), alpha, p(np.ascontiguousarray(x)), beta, p(bb), p(y))
    return y


def matvecT(A, alpha, x, beta, y):
    ii, jj, aa = csr_arrays(A)
    out = np.array(y, dtype=np.float64, copy=True)
    lib().orc_csr_matvecT(A.shape[0], A.shape[1], p(ii), p(jj), p(aa), alpha, p(np.ascontiguousarray(x)), beta, p(out))
    return out


def relax_l1jacobi(A, f, l1, w, u, all_zeros):
    ii, jj, aa = csr_arrays(A)
    u = np.array(u, dtype=np.float64, copy=True)
    vt = np.zeros_like(u)
    lib().orc_relax_l1jacobi(A.shape[0], p(ii), p(jj), p(aa), p(f), p(l1), w, p(u), p(vt), int(all_zeros))
    return u


def relax_cheby(A, f, ds, coefs, order, scale, u):
    ii, jj, aa = csr_arrays(A)
    u = np.array(u, dtype=np.float64, copy=True)
    n = u.size
    w = [np.zeros(n) for _ in range(4)]
    lib().orc_relax_cheby(n, p(ii), p(jj), p(aa), p(f), p(ds), p(np.ascontiguousarray(coefs)), order, scale,
                          p(u), p(w[0]), p(w[1]), p(w[2]), p(w[3]))
    return u


def relax_gs2(A, f, w, inner, u):
    ii, jj, aa = csr_arrays(A)
    u = np.array(u, dtype=np.float64, copy=True)
    vt = np.zeros_like(u)
    err = lib().orc_relax_gs2(u.size, p(ii), p(jj), p(aa), p(f), w, inner, p(u), p(vt))
    return u, err


class OracleAMG:
    """orc_amg built from the reference's hierarchy (hypre_b200.hypre.BoomerAMG after Setup)."""

    def __init__(self, amg):
        L = lib()
        self.keep = []
        nl = amg.num_levels
        prm = amg.params()
        self.h = C.c_void_p(L.orc_amg_create(nl))
        self.n = []
        for l in range(nl):
            Ad, _ = amg.A(l).blocks()
            n = Ad.shape[0]
            self.n.append(n)
            Ai, Aj, Aa = csr_arrays(Ad)
            if l < nl - 1:
                Pd, _ = amg.P(l).blocks()
                Pi, Pj, Pa = csr_arrays(Pd)
                nc = Pd.shape[1]
            else:
                Pi = Pj = Pa = None
                nc = 0
            l1 = amg.level_array("l1", l, n)
            ds = amg.level_array("cheby_ds", l, n)
            coefs = amg.level_array("cheby_coefs", l, prm["cheby_order"] + 1) if 16 in prm["relax_type"] else None
            cf = amg.cf_marker(l, n) if prm["relax_order"] == 1 else None
            self.keep.append((Ai, Aj, Aa, Pi, Pj, Pa, l1, ds, coefs, cf))
            L.orc_amg_set_level(self.h, l, n, p(Ai), p(Aj), p(Aa), nc, p(Pi), p(Pj), p(Pa), p(l1), p(ds), p(coefs),
                                amg.relax_weight(l))
            if cf is not None:
                L.orc_amg_set_level_cf(self.h, l, p(cf))
        rt = np.array(prm["relax_type"], dtype=np.int32)
        if rt[3] == 19:
            rt[3] = 9
        ns = np.array(prm["num_sweeps"], dtype=np.int32)
        L.orc_amg_set_cycle(self.h, p(rt), p(ns), prm["cycle_type"], prm["fcycle"], prm["cheby_order"],
                            prm["cheby_scale"])
        L.orc_amg_set_relax_order(self.h, prm["relax_order"])
        dense = amg.coarse_dense()
        if dense is not None:
            dense = np.ascontiguousarray(dense)
            L.orc_amg_set_coarse_dense(self.h, dense.shape[0], p(dense))
        # cycles per application when the solver is a preconditioner (tol 0; ij -precon_cycles)
        mi = C.c_int(1)
        amg.GetMaxIter(C.byref(mi))
        L.orc_amg_set_precond_cycles(self.h, mi.value)

    @classmethod
    def from_arrays(cls, levels, relax_type, num_sweeps, cycle_type, fcycle, cheby_order, cheby_scale, dense):
        """orc_amg from plain arrays: levels = [dict(A=csr, P=csr|None, l1=, ds=, coefs=, weight=)], e.g. a
        hierarchy gathered from N ranks of the reference (tests/test_oracle_multirank.py)"""
        self = cls.__new__(cls)
        L = lib()
        self.keep, self.n = [], []
        nl = len(levels)
        self.h = C.c_void_p(L.orc_amg_create(nl))
        for l, lv in enumerate(levels):
            Ai, Aj, Aa = csr_arrays(lv["A"])
            n = lv["A"].shape[0]
            self.n.append(n)
            if lv.get("P") is not None:
                Pi, Pj, Pa = csr_arrays(lv["P"]); nc = lv["P"].shape[1]
            else:
                Pi = Pj = Pa = None; nc = 0
            f64 = lambda a: None if a is None or len(a) == 0 else np.ascontiguousarray(a, dtype=np.float64)  # noqa: E731
            l1, ds, coefs = f64(lv.get("l1")), f64(lv.get("ds")), f64(lv.get("coefs"))
            self.keep.append((Ai, Aj, Aa, Pi, Pj, Pa, l1, ds, coefs))
            L.orc_amg_set_level(self.h, l, n, p(Ai), p(Aj), p(Aa), nc, p(Pi), p(Pj), p(Pa), p(l1), p(ds), p(coefs),
                                float(lv.get("weight", 1.0)))
        rt = np.array(relax_type, dtype=np.int32)
        if rt[3] == 19:
            rt[3] = 9
        ns = np.array(num_sweeps, dtype=np.int32)
        L.orc_amg_set_cycle(self.h, p(rt), p(ns), int(cycle_type), int(fcycle), int(cheby_order), int(cheby_scale))
        if dense is not None:
            dense = np.ascontiguousarray(dense, dtype=np.float64)
            self.keep.append(dense)
            L.orc_amg_set_coarse_dense(self.h, dense.shape[0], p(dense))
        return self

    def set_level_blocks(self, level, starts):
        """N-rank row partition of a level (two-stage GS uses the rank-local strict lower triangle)"""
        st = np.ascontiguousarray(starts, dtype=np.int32)
        lib().orc_amg_set_level_blocks(self.h, level, len(st) - 1, p(st))

    def cycle(self, f, u, all_zeros):
        u = np.array(u, dtype=np.float64, copy=True)
        lib().orc_amg_cycle(self.h, p(np.ascontiguousarray(f)), p(u), int(all_zeros))
        return u

    def pcg(self, b, x0, tol=1e-8, max_iter=100, two_norm=1):
        x = np.array(x0, dtype=np.float64, copy=True)
        norms = np.zeros(max_iter + 1)
        rel = np.zeros(max_iter + 1)
        conv = C.c_int()
        it = lib().orc_pcg(self.h, p(np.ascontiguousarray(b)), p(x), tol, max_iter, two_norm, p(norms), p(rel),
                           C.byref(conv))
        return x, it, rel[1:it + 1].copy(), conv.value

    def lgmres(self, b, x0, precond=0, tol=1e-8, a_tol=0.0, k_dim=5, aug_dim=2, approx_constant=1, min_iter=0,
               max_iter=100, skip_real_r_check=0):
        """orc_lgmres: (x, iterations, norms[0..iterations], final relative residual, converged)"""
        x = np.array(x0, dtype=np.float64, copy=True)
        norms = np.zeros(max_iter + 1)
        rel, conv = C.c_double(), C.c_int()
        it = lib().orc_lgmres(self.h, precond, p(np.ascontiguousarray(b)), p(x), tol, a_tol, k_dim, aug_dim,
                              approx_constant, min_iter, max_iter, p(norms), C.byref(rel), C.byref(conv))
        return x, it, norms[:it + 1].copy(), rel.value, conv.value

    def flexgmres(self, b, x0, **kw):
        """orc_flexgmres (preconditioned basis kept): same arguments and results as gmres"""
        return self.gmres(b, x0, _fn="orc_flexgmres", **kw)

    def cogmres(self, b, x0, **kw):
        """orc_cogmres (classical Gram-Schmidt): same arguments and results as gmres"""
        return self.gmres(b, x0, _fn="orc_cogmres", **kw)

    def gmres(self, b, x0, precond=0, tol=1e-8, a_tol=0.0, k_dim=5, min_iter=0, max_iter=100, skip_real_r_check=0,
              _fn="orc_gmres"):
        """orc_gmres: (x, iterations, norms[0..iterations], final relative residual, converged)"""
        x = np.array(x0, dtype=np.float64, copy=True)
        norms = np.zeros(max_iter + 1)
        rel, conv = C.c_double(), C.c_int()
        it = getattr(lib(), _fn)(self.h, precond, p(np.ascontiguousarray(b)), p(x), tol, a_tol, k_dim, min_iter, max_iter,
                             skip_real_r_check, p(norms), C.byref(rel), C.byref(conv))
        return x, it, norms[:it + 1].copy(), rel.value, conv.value

    def bicgstab(self, b, x0, precond=0, tol=1e-8, a_tol=0.0, min_iter=0, max_iter=100):
        """orc_bicgstab: (x, iterations (negative: breakdown), norms[0..iterations], final relative residual, converged)"""
        x = np.array(x0, dtype=np.float64, copy=True)
        norms = np.zeros(max_iter + 1)
        rel, conv = C.c_double(), C.c_int()
        it = lib().orc_bicgstab(self.h, precond, p(np.ascontiguousarray(b)), p(x), tol, a_tol, min_iter, max_iter,
                                p(norms), C.byref(rel), C.byref(conv))
        return x, it, norms[:max(it, 0) + 1].copy(), rel.value, conv.value

    def destroy(self):
        if self.h:
            lib().orc_amg_destroy(self.h)
            self.h = None


class krylov_options:
    """with krylov_options(cf_tol=0.7): ...   the oracle's Krylov solves inside read these (orc_krylov_set_options)"""

    def __init__(self, cf_tol=0.0, stop_crit=0):
        self.v = (float(cf_tol), int(stop_crit))

    def __enter__(self):
        lib().orc_krylov_set_options(*self.v)

    def __exit__(self, *a):
        lib().orc_krylov_set_options(0.0, 0)


def golden():
    return json.load(open(GOLDEN))


# the scaled-down versions of BASELINE.json's configs used across the tests
CASES = {
    "lap7_32_rlx18": dict(gen="laplacian", dims=(32, 32, 32), rlx=18),
    "lap7_100_rlx18": dict(gen="laplacian", dims=(100, 100, 100), rlx=18),
    "lap7_64_rlx16": dict(gen="laplacian", dims=(64, 64, 64), rlx=16),
    "lap27_48_rlx11": dict(gen="laplacian_27pt", dims=(48, 48, 48), rlx=11),
    "lap27_48_rlx12": dict(gen="laplacian_27pt", dims=(48, 48, 48), rlx=12),
    "rot_512_rlx18_agg1": dict(gen="rotate_7pt", dims=(512, 512, 45.0, 0.001), rlx=18, opts={"AggNumLevels": 1}),
    "lap7_24_rlx16": dict(gen="laplacian", dims=(24, 24, 24), rlx=16),
    "lap27_20_rlx11": dict(gen="laplacian_27pt", dims=(20, 20, 20), rlx=11),
    "lap27_20_rlx12": dict(gen="laplacian_27pt", dims=(20, 20, 20), rlx=12),
    # W- and F-cycles, two sweeps, weighted Jacobi (no goldens: checked against the reference in-process)
    "lap7_24_rlx18_W": dict(gen="laplacian", dims=(24, 24, 24), rlx=18, opts={"CycleType": 2}),
    "lap7_24_rlx16_F": dict(gen="laplacian", dims=(24, 24, 24), rlx=16, opts={"FCycle": 1}),
    "lap27_20_rlx18_ns2": dict(gen="laplacian_27pt", dims=(20, 20, 20), rlx=18, opts={"NumSweeps": 2}),
    "lap7_24_rlx7_w08": dict(gen="laplacian", dims=(24, 24, 24), rlx=7, opts={"RelaxWt": 0.8}),
    # CF-ordered relaxation (ij -CF 1 = HYPRE_BoomerAMGSetRelaxOrder(1)): C points then F points on the way down,
    # F then C on the way up (par_relax_interface.c:20-64)
    "lap7_24_rlx18_CF": dict(gen="laplacian", dims=(24, 24, 24), rlx=18, opts={"RelaxOrder": 1}),
    "lap27_20_rlx0_CF_w07": dict(gen="laplacian_27pt", dims=(20, 20, 20), rlx=0, opts={"RelaxOrder": 1, "RelaxWt": 0.7}),
    "lap7_24_rlx7_CF_W": dict(gen="laplacian", dims=(24, 24, 24), rlx=7, opts={"RelaxOrder": 1, "CycleType": 2, "RelaxWt": 0.9}),
    "lap27_20_rlx11_CF": dict(gen="laplacian_27pt", dims=(20, 20, 20), rlx=11, opts={"RelaxOrder": 1}),
    "lap7_24_rlx16_CF": dict(gen="laplacian", dims=(24, 24, 24), rlx=16, opts={"RelaxOrder": 1}),
    # relax_type 0: weighted point Jacobi (par_relax.c:320-335), five jobs of the reference's TEST_ij suites
    "lap7_24_rlx0_w08": dict(gen="laplacian", dims=(24, 24, 24), rlx=0, opts={"RelaxWt": 0.8}),
    "lap27_20_rlx0_w07_ns2": dict(gen="laplacian_27pt", dims=(20, 20, 20), rlx=0, opts={"RelaxWt": 0.7, "NumSweeps": 2}),
    # Chebyshev variant 1 (par_cheby.c:29-33: only the setup's coefficients differ), orders 3 and 4, unscaled
    # BoomerAMG as preconditioner with three cycles per application (ij -precon_cycles 3; TEST_ij/solvers.jobs:54)
    "lap7_24_rlx18_pc3": dict(gen="laplacian", dims=(24, 24, 24), rlx=18, opts={"MaxIter": 3}),
    "lap7_24_rlx16_pc2": dict(gen="laplacian", dims=(24, 24, 24), rlx=16, opts={"MaxIter": 2}),
    "lap7_24_rlx16_var1": dict(gen="laplacian", dims=(24, 24, 24), rlx=16, opts={"ChebyVariant": 1}),
    "lap7_24_rlx16_o4": dict(gen="laplacian", dims=(24, 24, 24), rlx=16, opts={"ChebyOrder": 4}),
    "lap7_24_rlx16_o3_noscale": dict(gen="laplacian", dims=(24, 24, 24), rlx=16, opts={"ChebyOrder": 3, "ChebyScale": 0}),
}

IJ_ARGS = {
    "lap7_32_rlx18": "-laplacian -n 32 32 32 -solver 1 -rlx 18",
    "lap7_100_rlx18": "-laplacian -n 100 100 100 -solver 1 -rlx 18",
    "lap7_64_rlx16": "-laplacian -n 64 64 64 -solver 1 -rlx 16",
    "lap27_48_rlx11": "-27pt -n 48 48 48 -solver 1 -rlx 11",
    "lap27_48_rlx12": "-27pt -n 48 48 48 -solver 1 -rlx 12",
    "rot_512_rlx18_agg1": "-rotate -n 512 512 -alpha 45 -eps 0.001 -solver 1 -rlx 18 -agg_nl 1",
    "lap7_24_rlx16": "-laplacian -n 24 24 24 -solver 1 -rlx 16",
    "lap27_20_rlx11": "-27pt -n 20 20 20 -solver 1 -rlx 11",
    "lap27_20_rlx12": "-27pt -n 20 20 20 -solver 1 -rlx 12",
}


def build_case(H, name, setup=True):
    c = CASES[name]
    A = getattr(H, c["gen"])(*c["dims"])
    b = A.vector(1.0)
    x = A.vector(0.0)
    amg = H.BoomerAMG(relax_type=c["rlx"], **c.get("opts", {}))
    pcg = H.ParCSRPCG(tol=1e-8, max_iter=100)
    pcg.SetPrecond(amg)
    if setup:
        pcg.Setup(A, b, x)
    return A, b, x, amg, pcg
