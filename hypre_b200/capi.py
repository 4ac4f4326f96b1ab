"""ctypes bindings of the C ABI declared in include/b200amg.h (libb200amg.so).

Fails loudly when the CUDA library is missing -- there is no CPU fallback.
"""
import ctypes as C
import os
import re

import numpy as np

from . import LIB_B200, ROOT

c_int_p = C.POINTER(C.c_int)
c_dbl_p = C.POINTER(C.c_double)
c_void_pp = C.POINTER(C.c_void_p)


class B200Error(RuntimeError):
    pass


class PcgParams(C.Structure):
    _fields_ = [("tol", C.c_double), ("a_tol", C.c_double), ("max_iter", C.c_int),
                ("two_norm", C.c_int), ("skip_break", C.c_int), ("diag_scale", C.c_int)]


class PcgResult(C.Structure):
    _fields_ = [("num_iterations", C.c_int), ("converged", C.c_int), ("error_flag", C.c_int),
                ("rel_residual_norm", C.c_double), ("bi_prod", C.c_double)]


class GmresParams(C.Structure):
    _fields_ = [("tol", C.c_double), ("a_tol", C.c_double), ("max_iter", C.c_int), ("min_iter", C.c_int),
                ("k_dim", C.c_int), ("skip_real_r_check", C.c_int), ("diag_scale", C.c_int), ("ortho", C.c_int),
                ("flex", C.c_int), ("cf_tol", C.c_double)]


class BicgstabParams(C.Structure):
    _fields_ = [("tol", C.c_double), ("a_tol", C.c_double), ("max_iter", C.c_int), ("min_iter", C.c_int),
                ("diag_scale", C.c_int), ("stop_crit", C.c_int), ("cf_tol", C.c_double)]


class LgmresParams(C.Structure):
    _fields_ = [("tol", C.c_double), ("a_tol", C.c_double), ("max_iter", C.c_int), ("min_iter", C.c_int),
                ("k_dim", C.c_int), ("aug_dim", C.c_int), ("approx_constant", C.c_int), ("diag_scale", C.c_int),
                ("cf_tol", C.c_double)]


class KrylovResult(C.Structure):
    _fields_ = [("num_iterations", C.c_int), ("converged", C.c_int), ("error_flag", C.c_int),
                ("rel_residual_norm", C.c_double), ("b_norm", C.c_double), ("real_residual_exit", C.c_int)]


def declared_symbols():
    """Every function name declared in include/b200amg.h."""
    text = open(os.path.join(ROOT, "include", "b200amg.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", text)))


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_B200):
        raise B200Error("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'`; "
                        "hypre_b200 has no CPU fallback" % LIB_B200)
    L = C.CDLL(LIB_B200, mode=C.RTLD_GLOBAL)
    vp, i, d, ll = C.c_void_p, C.c_int, C.c_double, C.c_longlong
    sig = {
        "b200_last_error": (C.c_char_p, []),
        "b200_version": (i, []),
        "b200_ctx_create": (i, [i, c_void_pp]),
        "b200_nccl_unique_id": (i, [C.c_char_p]),
        "b200_ctx_create_dist": (i, [i, i, i, C.c_char_p, c_void_pp]),
        "b200_ctx_destroy": (i, [vp]),
        "b200_ctx_sync": (i, [vp]),
        "b200_ctx_rank": (i, [vp, c_int_p, c_int_p]),
        "b200_ctx_launch_count": (ll, [vp]),
        "b200_timer_start": (i, [vp]),
        "b200_timer_stop": (i, [vp, C.POINTER(C.c_float)]),
        "b200_flush_l2": (i, [vp]),
        "b200_prof_report": (i, [vp, C.c_char_p, i]),
        "b200_dev_alloc": (i, [vp, ll, c_void_pp]),
        "b200_dev_free": (i, [vp, vp]),
        "b200_h2d": (i, [vp, vp, vp, ll]),
        "b200_d2h": (i, [vp, vp, vp, ll]),
        "b200_host_pin": (i, [vp, ll]),
        "b200_host_unpin": (i, [vp]),
        "b200_vec_fill": (i, [vp, i, d, vp]),
        "b200_vec_copy": (i, [vp, i, vp, vp]),
        "b200_vec_scale": (i, [vp, i, d, vp]),
        "b200_vec_axpy": (i, [vp, i, d, vp, vp]),
        "b200_vec_dot": (i, [vp, i, vp, vp, c_dbl_p]),
        "b200_vec_dot_noalloc": (i, [vp, i, vp, vp, vp]),
        "b200_parcsr_create": (i, [vp, i, i, vp, vp, vp, i, vp, vp, vp, i, vp, vp, vp, i, vp, vp, c_void_pp]),
        "b200_parcsr_destroy": (i, [vp]),
        "b200_parcsr_set_gs_blocks": (i, [vp, i, vp]),
        "b200_parcsr_info": (i, [vp, c_int_p, c_int_p, C.POINTER(ll), c_int_p, c_int_p]),
        "b200_set_layout_override": (i, [i, i]),
        "b200_set_compression": (i, [i, i]),
        "b200_set_kernel_options": (i, [i, i]),
        "b200_storage_decode_host": (i, [i, i, vp, vp, vp, vp, vp, c_int_p, c_int_p, C.POINTER(ll), C.POINTER(ll)]),
        "b200_parcsr_storage": (i, [vp, c_int_p, C.POINTER(ll), C.POINTER(ll), c_int_p]),
        "b200_rowpart_create": (i, [i, i, i, i, vp, vp, vp, vp, vp, c_void_pp]),
        "b200_rowpart_destroy": (i, [vp]),
        "b200_rowpart_sizes": (i, [vp] + [c_int_p] * 8),
        "b200_rowpart_int_array": (i, [vp, i, C.POINTER(c_int_p), c_int_p]),
        "b200_rowpart_real_array": (i, [vp, i, C.POINTER(c_dbl_p), c_int_p]),
        "b200_parcsr_create_from_rowpart": (i, [vp, vp, c_void_pp]),
        "b200_parcsr_matvec": (i, [vp, d, vp, d, vp, vp]),
        "b200_parcsr_matvecT": (i, [vp, d, vp, d, vp]),
        "b200_parcsr_matvec_host": (i, [vp, d, vp, d, vp, vp]),
        "b200_parcsr_matvecT_host": (i, [vp, d, vp, d, vp]),
        "b200_relax_l1jacobi": (i, [vp, vp, vp, d, vp, vp, i]),
        "b200_relax_cheby": (i, [vp, vp, vp, vp, i, i, vp, vp, vp, vp, i]),
        "b200_relax_2stage_gs": (i, [vp, vp, d, i, vp, vp, vp, i]),
        "b200_amg_create": (i, [vp, i, c_void_pp]),
        "b200_amg_destroy": (i, [vp]),
        "b200_amg_set_level": (i, [vp, i, vp, vp]),
        "b200_amg_set_cycle": (i, [vp, vp, vp, i, i]),
        "b200_amg_set_replicated_from": (i, [vp, i]),
        "b200_amg_set_level_relax": (i, [vp, i, d, vp]),
        "b200_amg_set_level_cheby": (i, [vp, i, vp, vp, i, i]),
        "b200_amg_set_coarse_dense": (i, [vp, i, vp, i, vp]),
        "b200_amg_set_relax_order": (i, [vp, i]),
        "b200_amg_set_level_cf_marker": (i, [vp, i, vp]),
        "b200_amg_finalize": (i, [vp]),
        "b200_amg_set_precond_cycles": (i, [vp, i]),
        "b200_amg_cycle": (i, [vp, vp, vp, i]),
        "b200_amg_cycle_bytes": (i, [vp, c_dbl_p]),
        "b200_amg_cycle_bytes_stored": (i, [vp, c_dbl_p]),
        "b200_cycle_sweep_counts": (i, [i, i, i, vp, vp]),
        "b200_cycle_exchange_counts": (i, [i, i, i, vp, vp, i, i, vp, vp, vp]),
        "b200_parcsr_halo_info": (i, [vp, c_int_p, c_int_p, c_int_p, c_int_p]),
        "b200_set_tail_limit": (i, [ll]),
        "b200_amg_tail_level": (i, [vp, c_int_p]),
        "b200_amg_level_matrix": (i, [vp, i, c_void_pp]),
        "b200_amg_level_interp": (i, [vp, i, c_void_pp]),
        "b200_amg_level_info": (i, [vp, i, c_int_p, C.POINTER(ll), C.POINTER(ll)]),
        "b200_amg_solve": (i, [vp, vp, vp, d, i, i, c_int_p, c_dbl_p, vp]),
        "b200_pcg_solve": (i, [vp, vp, vp, C.POINTER(PcgParams), vp, vp, C.POINTER(PcgResult), vp, vp]),
        "b200_pcg_solve_host": (i, [vp, vp, vp, C.POINTER(PcgParams), vp, vp, C.POINTER(PcgResult), vp, vp]),
        "b200_gmres_solve": (i, [vp, vp, vp, C.POINTER(GmresParams), vp, vp, C.POINTER(KrylovResult), vp]),
        "b200_gmres_solve_host": (i, [vp, vp, vp, C.POINTER(GmresParams), vp, vp, C.POINTER(KrylovResult), vp]),
        "b200_bicgstab_solve": (i, [vp, vp, vp, C.POINTER(BicgstabParams), vp, vp, C.POINTER(KrylovResult), vp]),
        "b200_bicgstab_solve_host": (i, [vp, vp, vp, C.POINTER(BicgstabParams), vp, vp, C.POINTER(KrylovResult), vp]),
        "b200_lgmres_solve": (i, [vp, vp, vp, C.POINTER(LgmresParams), vp, vp, C.POINTER(KrylovResult), vp]),
        "b200_lgmres_solve_host": (i, [vp, vp, vp, C.POINTER(LgmresParams), vp, vp, C.POINTER(KrylovResult), vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)      # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def check(status):
    if status != 0:
        raise B200Error("b200 error %d: %s" % (status, lib().b200_last_error().decode()))


def ptr(a):
    """host pointer of a numpy array (or None)."""
    if a is None:
        return None
    return a.ctypes.data_as(C.c_void_p)


class DeviceVector:
    """A device array of doubles owned through b200_dev_alloc."""

    def __init__(self, ctx, n, host=None):
        self.ctx, self.n = ctx, int(n)
        p = C.c_void_p()
        check(lib().b200_dev_alloc(ctx, self.n, C.byref(p)))
        self.ptr = p
        if host is not None:
            self.set(host)

    def set(self, host):
        h = np.ascontiguousarray(host, dtype=np.float64)
        assert h.size == self.n
        check(lib().b200_h2d(self.ctx, self.ptr, ptr(h), self.n))

    def get(self):
        h = np.empty(self.n, dtype=np.float64)
        check(lib().b200_d2h(self.ctx, ptr(h), self.ptr, self.n))
        return h

    def free(self):
        if self.ptr:
            lib().b200_dev_free(self.ctx, self.ptr)
            self.ptr = None


def create_context(device=0):
    ctx = C.c_void_p()
    check(lib().b200_ctx_create(device, C.byref(ctx)))
    return ctx


def parcsr_from_scipy(ctx, A):
    """single-rank b200_parcsr from a scipy CSR matrix (diag block only)."""
    A = A.tocsr()
    ii = np.ascontiguousarray(A.indptr, dtype=np.int32)
    jj = np.ascontiguousarray(A.indices, dtype=np.int32)
    aa = np.ascontiguousarray(A.data, dtype=np.float64)
    h = C.c_void_p()
    check(lib().b200_parcsr_create(ctx, A.shape[0], A.shape[1], ptr(ii), ptr(jj), ptr(aa),
                                   0, None, None, None, 0, None, None, None, 0, None, None, C.byref(h)))
    return h


class RowPart:
    """b200_rowpart: this rank's ParCSR blocks + halo plan of a replicated global CSR matrix
    (host only -- usable without a GPU)."""
    INT_NAMES = ["diag_i", "diag_j", "offd_i", "offd_j", "col_map_offd", "send_procs", "send_map_starts",
                 "send_map_elmts", "recv_procs", "recv_vec_starts"]

    def __init__(self, rank, nranks, A, row_starts, col_starts):
        A = A.tocsr()
        ii = np.ascontiguousarray(A.indptr, dtype=np.int32)
        jj = np.ascontiguousarray(A.indices, dtype=np.int32)
        aa = np.ascontiguousarray(A.data, dtype=np.float64)
        rs = np.ascontiguousarray(row_starts, dtype=np.int32)
        cs = np.ascontiguousarray(col_starts, dtype=np.int32)
        self.h = C.c_void_p()
        check(lib().b200_rowpart_create(rank, nranks, A.shape[0], A.shape[1], ptr(ii), ptr(jj), ptr(aa),
                                        ptr(rs), ptr(cs), C.byref(self.h)))
        for k, name in enumerate(self.INT_NAMES):
            p, n = c_int_p(), C.c_int()
            check(lib().b200_rowpart_int_array(self.h, k, C.byref(p), C.byref(n)))
            setattr(self, name, np.ctypeslib.as_array(p, (n.value,)).copy() if n.value else np.zeros(0, np.int32))
        for k, name in enumerate(["diag_a", "offd_a"]):
            p, n = c_dbl_p(), C.c_int()
            check(lib().b200_rowpart_real_array(self.h, k, C.byref(p), C.byref(n)))
            setattr(self, name, np.ctypeslib.as_array(p, (n.value,)).copy() if n.value else np.zeros(0))
        s = [C.c_int() for _ in range(8)]
        check(lib().b200_rowpart_sizes(self.h, *[C.byref(v) for v in s]))
        (self.n_rows, self.n_cols_diag, self.n_cols_offd, self.nnz_diag, self.nnz_offd, self.num_sends,
         self.n_send_elmts, self.num_recvs) = [v.value for v in s]

    def blocks(self):
        import scipy.sparse as sp
        diag = sp.csr_matrix((self.diag_a, self.diag_j, self.diag_i), shape=(self.n_rows, self.n_cols_diag))
        offd = sp.csr_matrix((self.offd_a, self.offd_j, self.offd_i), shape=(self.n_rows, self.n_cols_offd))
        return diag, offd

    def device(self, ctx):
        h = C.c_void_p()
        check(lib().b200_parcsr_create_from_rowpart(ctx, self.h, C.byref(h)))
        return h

    def destroy(self):
        if self.h:
            lib().b200_rowpart_destroy(self.h)
            self.h = None


def even_starts(n, nranks):
    """hypre_GeneratePartitioning: the first n % nranks ranks get one extra row."""
    size, rest = divmod(n, nranks)
    st = np.zeros(nranks + 1, dtype=np.int32)
    for r in range(nranks):
        st[r + 1] = st[r] + size + (1 if r < rest else 0)
    return st
