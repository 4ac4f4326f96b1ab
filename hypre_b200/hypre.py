"""Python mirror of the part of the HYPRE C API that src/test/ij.c uses for
`-solver 1` (AMG-PCG), `-solver 0` (AMG) and `-solver -1` (SpMV), bound with ctypes
to libHYPRE_b200.so (the B200 binding layered over the reference's host library).

Names and argument meaning follow the reference (HYPRE_parcsr_ls.h / HYPRE_krylov.h):
``BoomerAMG.SetRelaxType(18)`` is ``HYPRE_BoomerAMGSetRelaxType(solver, 18)`` and so on.
``set_device_path(False)`` routes the interposed entry points to the reference's own CPU
implementation in the same process (test / cpu_baseline use only).
"""
import ctypes as C
import os

import numpy as np

from . import LIB_BRIDGE, LIB_BRIDGE_MP, LIB_HYPRE_B200, LIB_HYPRE_B200_MP, LIB_HYPRE_REF, LIB_REF_BRIDGE, capi

_lib = None
# Reference-only mode (bench.py --impl reference, the cpu_baseline leg): this process loads the oracle's separately
# built reference library (oracle/_ref/libHYPRE_ref.so) and NOTHING of the product -- no libb200amg.so, no binding,
# no product host library -- so that what is timed is the unmodified reference alone.  Device entry points do not
# exist in this mode.
REFERENCE_ONLY = os.environ.get("HYPRE_B200_REFERENCE_ONLY", "0") == "1"


class _Libs:
    """symbol lookup over the hypre library of this process and the accessor bridge (tests / bench plumbing)"""

    def __init__(self, main, bridge):
        self._main, self._bridge = main, bridge

    def __getattr__(self, name):
        try:
            f = getattr(self._bridge if name.startswith("hb_") else self._main, name)
        except AttributeError:
            f = getattr(self._main if name.startswith("hb_") else self._bridge, name)
        setattr(self, name, f)
        return f

COMM_WORLD = 0   # hypre_MPI_COMM_WORLD of the sequential stubs (src/utilities/mpistubs.h) and of hypre_b200/host/mpishim
# Native multi-rank mode: this process is one of N ranks (one per GPU), the
# reference is its MPI build over hypre_b200/host/mpishim and hypre's own distributed setup runs.  Ranks started by
# torchrun meet through SHIM_DIR (hypre_b200/host/mpishim/mpishim.c); under mpirun_shim nothing else is needed.
NATIVE = os.environ.get("HYPRE_B200_NATIVE", "0") == "1"


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if REFERENCE_ONLY:
        if NATIVE:
            raise capi.B200Error("reference-only mode is single-rank")
        path = LIB_HYPRE_REF
    else:
        capi.lib()   # libb200amg.so first: fail loudly if the CUDA library is missing
        path = LIB_HYPRE_B200_MP if NATIVE else LIB_HYPRE_B200
    if not os.path.exists(path):
        raise capi.B200Error("%s is missing (built from hypre_b200/host or oracle/ with the reference sources)" % path)
    main = C.CDLL(path, mode=C.RTLD_GLOBAL)
    bridge_path = LIB_REF_BRIDGE if REFERENCE_ONLY else LIB_BRIDGE_MP if NATIVE else LIB_BRIDGE
    if not os.path.exists(bridge_path):
        raise capi.B200Error("%s is missing (built from hypre_b200/host)" % bridge_path)
    L = _Libs(main, C.CDLL(bridge_path, mode=C.RTLD_GLOBAL))
    if NATIVE:
        L.MPI_Init(None, None)
    vp, i, d = C.c_void_p, C.c_int, C.c_double
    for name in ("GenerateLaplacian", "GenerateLaplacian27pt", "GenerateRotate7pt", "hb_vector_create",
                 "hb_vector_create_cols", "hb_amg_A", "hb_amg_P", "hb_amg_F", "hb_amg_U", "hb_amg_Vtemp",
                 "hb_amg_Ztemp", "hb_amg_Ptemp", "hb_amg_Rtemp", "hb_pcg_precond_data") + \
            (() if REFERENCE_ONLY else ("hypre_b200_context", "hypre_b200_amg_mirror", "hypre_b200_matrix_mirror")):
        getattr(L, name).restype = vp
    for name in ("hb_vector_data", "hb_amg_l1", "hb_amg_cheby_ds", "hb_amg_cheby_coefs",
                 "hb_amg_coarse_dense", "hb_pcg_norms", "hb_pcg_rel_norms", "hb_gmres_norms", "hb_cogmres_norms", "hb_flexgmres_norms", "hb_lgmres_norms", "hb_bicgstab_norms"):
        getattr(L, name).restype = C.POINTER(d)
    L.hb_amg_relax_weight.restype = d
    L.hb_amg_cf_marker.restype = C.POINTER(C.c_int)
    L.hb_amg_cf_marker.argtypes = [vp, i]
    L.hb_amg_omega.restype = d
    L.hb_matrix_global_rows.restype = C.c_longlong
    L.GenerateLaplacian.argtypes = [i, i, i, i, i, i, i, i, i, i, vp]
    L.GenerateLaplacian27pt.argtypes = [i, i, i, i, i, i, i, i, i, i, vp]
    L.GenerateRotate7pt.argtypes = [i, i, i, i, i, i, i, d, d]
    L.hb_vector_create.argtypes = [vp, d]
    L.hb_vector_create_cols.argtypes = [vp, d]
    for name in ("hb_vector_destroy", "hb_vector_data", "hb_vector_size", "hb_amg_num_levels", "hb_amg_Vtemp",
                 "hb_amg_Ztemp", "hb_amg_Ptemp", "hb_amg_Rtemp", "hb_pcg_norms", "hb_pcg_rel_norms",
                 "hb_gmres_norms", "hb_cogmres_norms", "hb_bicgstab_norms", "hb_bicgstab_converged", "HYPRE_ParCSRGMRESDestroy",
                 "HYPRE_ParCSRCOGMRESDestroy", "hb_flexgmres_norms", "HYPRE_ParCSRFlexGMRESDestroy", "hb_lgmres_norms", "HYPRE_ParCSRLGMRESDestroy", "HYPRE_ParCSRBiCGSTABDestroy",
                 "hb_pcg_precond_data", "hb_matrix_global_rows", "HYPRE_ParCSRMatrixDestroy", "HYPRE_BoomerAMGDestroy",
                 "HYPRE_ParCSRPCGDestroy") + (() if REFERENCE_ONLY else ("hypre_b200_amg_mirror", "hypre_b200_matrix_mirror")):
        getattr(L, name).argtypes = [vp]
    for name in ("hb_amg_A", "hb_amg_P", "hb_amg_F", "hb_amg_U", "hb_amg_l1", "hb_amg_cheby_ds",
                 "hb_amg_cheby_coefs", "hb_amg_relax_weight", "hb_amg_omega", "hb_vector_set_all_zeros"):
        getattr(L, name).argtypes = [vp, i]
    L.hb_amg_params.argtypes = [vp, vp]
    L.hb_amg_coarse_dense.argtypes = [vp, C.POINTER(i)]
    L.hb_matrix_blocks.argtypes = [vp] + [vp] * 9
    if not REFERENCE_ONLY:
        L.hypre_b200_set_enabled.argtypes = [i]
    L.HYPRE_ParCSRMatrixMatvec.argtypes = [d, vp, vp, d, vp]
    L.HYPRE_ParCSRMatrixMatvecT.argtypes = [d, vp, vp, d, vp]
    L.hypre_ParCSRMatrixMatvecOutOfPlace.argtypes = [d, vp, vp, d, vp, vp]
    L.hypre_ParCSRMatrixMatvecT.argtypes = [d, vp, vp, d, vp]
    L.hypre_BoomerAMGRelax.argtypes = [vp, vp, vp, i, i, d, d, vp, vp, vp, vp]
    L.hypre_ParCSRRelax_Cheby_Solve.argtypes = [vp, vp, vp, vp, i, i, i, vp, vp, vp, vp, vp]
    L.hypre_BoomerAMGCycle.argtypes = [vp, vp, vp]
    L.HYPRE_BoomerAMGSolve.argtypes = [vp, vp, vp, vp]
    L.HYPRE_BoomerAMGSetup.argtypes = [vp, vp, vp, vp]
    L.HYPRE_Initialize()
    _lib = L
    return L


def world():
    """(rank, size) of hypre's world communicator; (0, 1) outside the native multi-rank mode"""
    if not NATIVE:
        return 0, 1
    r, n = C.c_int(), C.c_int()
    lib().MPI_Comm_rank(COMM_WORLD, C.byref(r))
    lib().MPI_Comm_size(COMM_WORLD, C.byref(n))
    return r.value, n.value


def process_grid(size):
    """P x Q x R with P >= Q >= R, as balanced as the factors of `size` allow (8 -> 2 2 2, 4 -> 2 2 1)"""
    best = (size, 1, 1)
    for r in range(1, size + 1):
        for q in range(r, size + 1):
            if size % (r * q) == 0 and size // (r * q) >= q:
                cand = (size // (r * q), q, r)
                if cand[0] - cand[2] < best[0] - best[2]:
                    best = cand
    return best


def _grid_args(grid):
    """(P, Q, R, p, q, r) for the generators: this rank's box of the process grid (ij.c:9620-9626)"""
    rank, size = world()
    if size == 1:
        return 1, 1, 1, 0, 0, 0
    P, Q, R = grid if grid else (size, 1, 1)
    if P * Q * R != size:
        raise ValueError("process grid %s does not match %d ranks" % ((P, Q, R), size))
    p = rank % P
    q = ((rank - p) // P) % Q
    r = (rank - p - P * q) // (P * Q)
    return P, Q, R, p, q, r


def set_device_path(enabled):
    """False: interposed entry points call the host library's own (CPU) implementation.  Tests only."""
    if REFERENCE_ONLY:
        if enabled:
            raise capi.B200Error("reference-only mode: the device path is not loaded in this process")
        lib()
        return
    lib().hypre_b200_set_enabled(1 if enabled else 0)


def error_flag():
    return lib().hb_error_flag()


def clear_errors():
    lib().hb_clear_errors()


# ------------------------------------------------------------------ matrices / vectors
class ParCSRMatrix:
    def __init__(self, handle, own=True):
        self.h, self.own = C.c_void_p(handle), own

    def blocks(self):
        """(diag, offd) as scipy CSR views of hypre's own arrays (copied)."""
        import scipy.sparse as sp
        L = lib()
        n, nc, no = C.c_int(), C.c_int(), C.c_int()
        di, dj, oi, oj = (C.POINTER(C.c_int)() for _ in range(4))
        da, oa = C.POINTER(C.c_double)(), C.POINTER(C.c_double)()
        L.hb_matrix_blocks(self.h, C.byref(n), C.byref(nc), C.byref(di), C.byref(dj), C.byref(da),
                           C.byref(no), C.byref(oi), C.byref(oj), C.byref(oa))
        n, nc, no = n.value, nc.value, no.value
        ii = np.ctypeslib.as_array(di, (n + 1,)).copy() if n else np.zeros(1, np.int32)
        nnz = int(ii[-1])
        jj = np.ctypeslib.as_array(dj, (max(nnz, 1),))[:nnz].copy()
        aa = np.ctypeslib.as_array(da, (max(nnz, 1),))[:nnz].copy()
        diag = sp.csr_matrix((aa, jj, ii), shape=(n, nc))
        offd = sp.csr_matrix((n, no))
        if no > 0:
            oii = np.ctypeslib.as_array(oi, (n + 1,)).copy()
            onz = int(oii[-1])
            offd = sp.csr_matrix((np.ctypeslib.as_array(oa, (max(onz, 1),))[:onz].copy(),
                                  np.ctypeslib.as_array(oj, (max(onz, 1),))[:onz].copy(), oii), shape=(n, no))
        return diag, offd

    @property
    def global_rows(self):
        return int(lib().hb_matrix_global_rows(self.h))

    def vector(self, value=0.0, cols=False):
        fn = lib().hb_vector_create_cols if cols else lib().hb_vector_create
        return ParVector(fn(self.h, float(value)))

    def device_mirror(self):
        m = lib().hypre_b200_matrix_mirror(self.h)
        if not m:
            raise capi.B200Error("matrix upload failed: " + capi.lib().b200_last_error().decode())
        return C.c_void_p(m)

    def destroy(self):
        if self.own and self.h:
            lib().HYPRE_ParCSRMatrixDestroy(self.h)
            self.h = None


class ParVector:
    def __init__(self, handle, own=True):
        self.h, self.own = C.c_void_p(handle), own

    @property
    def size(self):
        return lib().hb_vector_size(self.h)

    def numpy(self):
        """numpy view of hypre's own host array (no copy)."""
        n = self.size
        if n == 0:
            return np.zeros(0)
        return np.ctypeslib.as_array(lib().hb_vector_data(self.h), (n,))

    # ---- device-resident form (hypre_glue.c "device-resident vectors")
    def bind_device(self, dev_ptr):
        """state that this vector's local part lives in the device array `dev_ptr` (capi.DeviceVector.ptr):
        the Krylov bindings then solve on it in place; None removes the binding"""
        L = lib()
        L.hypre_b200_vector_bind_device.argtypes = [C.c_void_p, C.c_void_p]
        return L.hypre_b200_vector_bind_device(self.h, dev_ptr)

    def bound_device(self):
        L = lib()
        L.hypre_b200_vector_device.argtypes = [C.c_void_p]
        L.hypre_b200_vector_device.restype = C.c_void_p
        return L.hypre_b200_vector_device(self.h)

    def sync_to_host(self, first=0, n=-1):
        L = lib()
        L.hypre_b200_vector_sync_to_host.argtypes = [C.c_void_p, C.c_int, C.c_int]
        return L.hypre_b200_vector_sync_to_host(self.h, first, n)

    def sync_to_device(self, first=0, n=-1):
        L = lib()
        L.hypre_b200_vector_sync_to_device.argtypes = [C.c_void_p, C.c_int, C.c_int]
        return L.hypre_b200_vector_sync_to_device(self.h, first, n)

    def set_all_zeros_flag(self, flag):
        lib().hb_vector_set_all_zeros(self.h, int(flag))

    def destroy(self):
        if self.own and self.h:
            lib().hb_vector_destroy(self.h)
            self.h = None


def laplacian(nx, ny, nz, c=(1.0, 1.0, 1.0), grid=None):
    """ij -laplacian -n nx ny nz [-c cx cy cz]   (BuildParLaplacian, src/test/ij.c:9559-9790)."""
    cx, cy, cz = c
    v = np.zeros(4)
    v[1], v[2], v[3] = -cx, -cy, -cz
    v[0] = (2.0 * cx if nx > 1 else 0.0) + (2.0 * cy if ny > 1 else 0.0) + (2.0 * cz if nz > 1 else 0.0)
    return ParCSRMatrix(lib().GenerateLaplacian(COMM_WORLD, nx, ny, nz, *_grid_args(grid), capi.ptr(v)))


def laplacian_27pt(nx, ny, nz, grid=None):
    """ij -27pt -n nx ny nz   (BuildParLaplacian27pt, src/test/ij.c:10984-10996)."""
    v = np.array([26.0, -1.0])
    if nx == 1 or ny == 1 or nz == 1:
        v[0] = 8.0
    if nx * ny == 1 or nx * nz == 1 or ny * nz == 1:
        v[0] = 2.0
    return ParCSRMatrix(lib().GenerateLaplacian27pt(COMM_WORLD, nx, ny, nz, *_grid_args(grid), capi.ptr(v)))


def rotate_7pt(nx, ny, alpha, eps):
    """ij -rotate -n nx ny -alpha a -eps e   (BuildParRotate7pt, src/test/ij.c:11184-11240)."""
    return ParCSRMatrix(lib().GenerateRotate7pt(COMM_WORLD, nx, ny, 1, 1, 0, 0, float(alpha), float(eps)))


# ------------------------------------------------------------------ solvers
class _Solver:
    _prefix = ""

    def __getattr__(self, name):
        # solver.SetXxx(args) -> HYPRE_<prefix>SetXxx(handle, args)
        if name.startswith(("Set", "Get")):
            fn = getattr(lib(), self._prefix + name)

            def call(*args):
                conv = [C.c_double(a) if isinstance(a, float) else a for a in args]
                return fn(self.h, *conv)
            return call
        raise AttributeError(name)


class BoomerAMG(_Solver):
    """HYPRE_BoomerAMG* (src/parcsr_ls/HYPRE_parcsr_amg.c) with the defaults ij.c passes."""
    _prefix = "HYPRE_BoomerAMG"

    def __init__(self, relax_type=18, as_precond=True, **opts):
        self.h = C.c_void_p()
        lib().HYPRE_BoomerAMGCreate(C.byref(self.h))
        # ij.c defaults (src/test/ij.c:248-302): HMIS, ext+i, Pmax 4, relax_order 0, max coarse 9
        self.SetCoarsenType(10)
        self.SetInterpType(6)
        self.SetPMaxElmts(4)
        self.SetRelaxOrder(0)
        self.SetMaxCoarseSize(9)
        self.SetMinCoarseSize(0)
        self.SetMaxRowSum(1.0)          # ij.c: max_row_sum = 1. (library default 0.9)
        self.SetStrongThreshold(0.25)
        self.SetMaxLevels(25)
        self.SetMeasureType(0)
        self.SetNumSweeps(1)
        self.SetTruncFactor(0.0)
        self.SetAggNumLevels(0)
        self.SetPrintLevel(0)
        self.SetChebyOrder(2)
        self.SetChebyEigEst(10)
        self.SetChebyVariant(0)
        self.SetChebyScale(1)
        self.SetChebyFraction(0.3)
        if relax_type is not None and relax_type >= 0:
            self.SetRelaxType(relax_type)
        if as_precond:
            self.SetTol(0.0)       # ij.c:5037 pc_tol = 0
            self.SetMaxIter(1)     # ij.c:5055 precon_cycles = 1
        for k, v in opts.items():
            getattr(self, "Set" + k)(v)

    def Setup(self, A, b, x):
        return lib().HYPRE_BoomerAMGSetup(self.h, A.h, b.h, x.h)

    def Solve(self, A, b, x):
        return lib().HYPRE_BoomerAMGSolve(self.h, A.h, b.h, x.h)

    def device_mirror(self):
        m = lib().hypre_b200_amg_mirror(self.h)
        if not m:
            raise capi.B200Error("hierarchy upload failed: " + capi.lib().b200_last_error().decode())
        return C.c_void_p(m)

    # hierarchy access (reference's host objects)
    @property
    def num_levels(self):
        return lib().hb_amg_num_levels(self.h)

    def A(self, level):
        return ParCSRMatrix(lib().hb_amg_A(self.h, level), own=False)

    def P(self, level):
        return ParCSRMatrix(lib().hb_amg_P(self.h, level), own=False)

    def params(self):
        out = np.zeros(14, dtype=np.int32)
        lib().hb_amg_params(self.h, capi.ptr(out))
        return {"relax_type": out[0:4].tolist(), "num_sweeps": out[4:8].tolist(), "cycle_type": int(out[8]),
                "fcycle": int(out[9]), "cheby_order": int(out[10]), "cheby_scale": int(out[11]),
                "cheby_variant": int(out[12]), "relax_order": int(out[13])}

    def level_array(self, which, level, n):
        p = getattr(lib(), "hb_amg_" + which)(self.h, level)
        return np.ctypeslib.as_array(p, (n,)).copy() if p else None

    def cf_marker(self, level, n):
        """CF_marker_array[level] (1 = C point, -1 = F point), None on the coarsest level"""
        p = lib().hb_amg_cf_marker(self.h, level)
        return np.ctypeslib.as_array(p, (n,)).copy().astype(np.int32) if p else None

    def relax_weight(self, level):
        return lib().hb_amg_relax_weight(self.h, level)

    def coarse_dense(self):
        n = C.c_int()
        p = lib().hb_amg_coarse_dense(self.h, C.byref(n))
        if not p or n.value == 0:
            return None
        return np.ctypeslib.as_array(p, (n.value * n.value,)).copy().reshape(n.value, n.value)

    def destroy(self):
        if self.h:
            lib().HYPRE_BoomerAMGDestroy(self.h)
            self.h = None


class ParCSRPCG(_Solver):
    """HYPRE_ParCSRPCG* / HYPRE_PCG* (src/parcsr_ls/HYPRE_parcsr_pcg.c, src/krylov/HYPRE_pcg.c)."""
    _prefix = "HYPRE_PCG"

    def __init__(self, tol=1e-8, max_iter=100, two_norm=1, print_level=0, logging=1):
        self.h = C.c_void_p()
        lib().HYPRE_ParCSRPCGCreate(COMM_WORLD, C.byref(self.h))
        self.SetMaxIter(max_iter)
        self.SetTol(float(tol))
        self.SetTwoNorm(two_norm)
        self.SetPrintLevel(print_level)
        self.SetLogging(logging)
        self.max_iter = max_iter
        self.precond = None

    def SetPrecond(self, amg):
        L = lib()
        self.precond = amg
        return L.HYPRE_PCGSetPrecond(self.h, L.HYPRE_BoomerAMGSolve, L.HYPRE_BoomerAMGSetup, amg.h)

    def SetDiagScalePrecond(self):
        L = lib()
        return L.HYPRE_PCGSetPrecond(self.h, L.HYPRE_ParCSRDiagScale, L.HYPRE_ParCSRDiagScaleSetup, None)

    def Setup(self, A, b, x):
        return lib().HYPRE_ParCSRPCGSetup(self.h, A.h, b.h, x.h)

    def Solve(self, A, b, x):
        return lib().HYPRE_ParCSRPCGSolve(self.h, A.h, b.h, x.h)

    @property
    def num_iterations(self):
        n = C.c_int()
        lib().HYPRE_PCGGetNumIterations(self.h, C.byref(n))
        return n.value

    @property
    def final_rel_norm(self):
        v = C.c_double()
        lib().HYPRE_PCGGetFinalRelativeResidualNorm(self.h, C.byref(v))
        return v.value

    def rel_norms(self):
        """per-iteration ||r||/||b|| (pcg.c:767-771), entries 1..num_iterations."""
        p = lib().hb_pcg_rel_norms(self.h)
        n = self.num_iterations
        return np.ctypeslib.as_array(p, (self.max_iter + 1,))[1:n + 1].copy()

    def destroy(self):
        if self.h:
            lib().HYPRE_ParCSRPCGDestroy(self.h)
            self.h = None


class _ParCSRKrylov(_Solver):
    """shared shape of HYPRE_ParCSRGMRES* / HYPRE_ParCSRBiCGSTAB* (src/parcsr_ls/HYPRE_parcsr_gmres.c,
    HYPRE_parcsr_bicgstab.c; setters in src/krylov/HYPRE_gmres.c, HYPRE_bicgstab.c)"""
    _prefix = ""
    _par = ""
    _norms = ""

    def __init__(self, tol=1e-8, max_iter=100, print_level=0, logging=1):
        self.h = C.c_void_p()
        getattr(lib(), self._par + "Create")(COMM_WORLD, C.byref(self.h))
        self.SetMaxIter(max_iter)
        self.SetTol(float(tol))
        self.SetPrintLevel(print_level)
        self.SetLogging(logging)
        self.max_iter = max_iter
        self.precond = None

    def SetPrecond(self, amg):
        L = lib()
        self.precond = amg
        return getattr(L, self._prefix + "SetPrecond")(self.h, L.HYPRE_BoomerAMGSolve, L.HYPRE_BoomerAMGSetup, amg.h)

    def SetDiagScalePrecond(self):
        L = lib()
        return getattr(L, self._prefix + "SetPrecond")(self.h, L.HYPRE_ParCSRDiagScale, L.HYPRE_ParCSRDiagScaleSetup, None)

    def Setup(self, A, b, x):
        return getattr(lib(), self._par + "Setup")(self.h, A.h, b.h, x.h)

    def Solve(self, A, b, x):
        return getattr(lib(), self._par + "Solve")(self.h, A.h, b.h, x.h)

    @property
    def num_iterations(self):
        n = C.c_int()
        getattr(lib(), self._prefix + "GetNumIterations")(self.h, C.byref(n))
        return n.value

    @property
    def final_rel_norm(self):
        v = C.c_double()
        getattr(lib(), self._prefix + "GetFinalRelativeResidualNorm")(self.h, C.byref(v))
        return v.value

    @property
    def converged(self):
        n = C.c_int()
        getattr(lib(), self._prefix + "GetConverged")(self.h, C.byref(n))
        return n.value

    def norms(self):
        """residual norms the solver logged: entries 0..num_iterations"""
        p = getattr(lib(), self._norms)(self.h)
        return np.ctypeslib.as_array(p, (self.max_iter + 1,))[:self.num_iterations + 1].copy()

    def destroy(self):
        if self.h:
            getattr(lib(), self._par + "Destroy")(self.h)
            self.h = None


class ParCSRGMRES(_ParCSRKrylov):
    """HYPRE_ParCSRGMRES* / HYPRE_GMRES*: ij -solver 3 (AMG) / 4 (diagonal scaling), k_dim 5 (ij.c:1731)."""
    _prefix = "HYPRE_GMRES"
    _par = "HYPRE_ParCSRGMRES"
    _norms = "hb_gmres_norms"

    def __init__(self, k_dim=5, **kw):
        super().__init__(**kw)
        self.SetKDim(k_dim)


class ParCSRCOGMRES(_ParCSRKrylov):
    """HYPRE_ParCSRCOGMRES* / HYPRE_COGMRES*: ij -solver 16 (AMG) / 17 (diagonal scaling); cgs 1, unroll 0 (ij.c defaults)."""
    _prefix = "HYPRE_COGMRES"
    _par = "HYPRE_ParCSRCOGMRES"
    _norms = "hb_cogmres_norms"

    def __init__(self, k_dim=5, **kw):
        super().__init__(**kw)
        self.SetKDim(k_dim)


class ParCSRFlexGMRES(_ParCSRKrylov):
    """HYPRE_ParCSRFlexGMRES* / HYPRE_FlexGMRES*: ij -solver 61 (AMG) / 60 (diagonal scaling), default modify_pc."""
    _prefix = "HYPRE_FlexGMRES"
    _par = "HYPRE_ParCSRFlexGMRES"
    _norms = "hb_flexgmres_norms"

    def __init__(self, k_dim=5, **kw):
        super().__init__(**kw)
        self.SetKDim(k_dim)


class ParCSRLGMRES(_ParCSRKrylov):
    """HYPRE_ParCSRLGMRES* / HYPRE_LGMRES*: ij -solver 51 (AMG) / 50 (diagonal scaling); aug_dim 2 (ij.c / lgmres.c:104)."""
    _prefix = "HYPRE_LGMRES"
    _par = "HYPRE_ParCSRLGMRES"
    _norms = "hb_lgmres_norms"

    def __init__(self, k_dim=5, aug_dim=2, **kw):
        super().__init__(**kw)
        self.SetKDim(k_dim)
        self.SetAugDim(aug_dim)


class ParCSRBiCGSTAB(_ParCSRKrylov):
    """HYPRE_ParCSRBiCGSTAB* / HYPRE_BiCGSTAB*: ij -solver 9 (AMG) / 10 (diagonal scaling)."""
    _prefix = "HYPRE_BiCGSTAB"
    _par = "HYPRE_ParCSRBiCGSTAB"
    _norms = "hb_bicgstab_norms"

    @property
    def converged(self):
        # no HYPRE_BiCGSTABGetConverged in the reference's API: read hypre_BiCGSTABData.converged
        return lib().hb_bicgstab_converged(self.h)
