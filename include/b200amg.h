/* b200amg.h -- C ABI of the B200 (sm_100a) BoomerAMG-PCG solve phase.
 *
 * Plain C types only: opaque handles, host pointers (suffix _h / "host"),
 * DEVICE pointers (double* documented as "dev"), ints and doubles.  Every
 * function returns B200_OK (0) or a negative B200_ERR_* code; the text of the
 * last error is available from b200_last_error().  There is NO CPU fallback in
 * this library: without a CUDA device every entry point that computes fails
 * with B200_ERR_CUDA.
 *
 * Each entry point names the reference (hypre 2.32.0, paths under src/) routine
 * whose solve-phase work it replaces.  Index type = int (HYPRE_Int), values =
 * double (HYPRE_Real), see src/utilities/HYPRE_utilities.h:35-61,105,131.
 */
#ifndef B200AMG_H
#define B200AMG_H

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK               0
#define B200_ERR_CUDA        -1
#define B200_ERR_ARG         -2
#define B200_ERR_NCCL        -3
#define B200_ERR_UNSUPPORTED -4
#define B200_ERR_NUMERIC     -5   /* zero diagonal, NaN/Inf input, ... */

typedef struct b200_ctx    b200_ctx;     /* device + streams + (optional) NCCL communicator */
typedef struct b200_parcsr b200_parcsr;  /* device mirror of hypre_ParCSRMatrix             */
typedef struct b200_amg    b200_amg;     /* device mirror of hypre_ParAMGData (solve part)  */

const char *b200_last_error(void);
int  b200_version(void);

/* ------------------------------------------------------------------ context
 * One context per process = one GPU (the reference binds one MPI rank to one
 * device: src/utilities/device_utils.c:3008-3068; compute stream =
 * hypre_HandleComputeStream, src/utilities/device_utils.h:252-269).
 * b200_ctx_create_dist joins an NCCL communicator (replaces the MPI_Comm of
 * hypre_ParCSRMatrixComm for the solve phase); uid = 128-byte ncclUniqueId
 * produced by b200_nccl_unique_id on rank 0 and broadcast by the caller. */
int b200_ctx_create(int device, b200_ctx **ctx);
int b200_nccl_unique_id(char uid[128]);
int b200_ctx_create_dist(int device, int rank, int nranks, const char uid[128], b200_ctx **ctx);
int b200_ctx_destroy(b200_ctx *ctx);
int b200_ctx_sync(b200_ctx *ctx);
int b200_ctx_rank(const b200_ctx *ctx, int *rank, int *nranks);
/* number of kernels this library launched on ctx since creation */
long long b200_ctx_launch_count(const b200_ctx *ctx);
/* CUDA-event timing on the library's own compute stream (ms) */
int b200_timer_start(b200_ctx *ctx);
int b200_timer_stop(b200_ctx *ctx, float *ms);
/* B200_PROFILE=1 in the environment brackets every launch with CUDA events; this returns the
 * per-kernel table (launches, total ms, average us) accumulated since the last call */
int b200_prof_report(b200_ctx *ctx, char *buf, int len);
/* write `bytes` of scratch (evicts L2 between timed iterations) */
int b200_flush_l2(b200_ctx *ctx);

/* ------------------------------------------------------------ device memory
 * hypre_TAlloc/hypre_TFree/hypre_TMemcpy(…, HYPRE_MEMORY_DEVICE)
 * (src/utilities/memory.c:817-842) for vectors of doubles. */
int b200_dev_alloc(b200_ctx *ctx, long long n, double **dev);
int b200_dev_free(b200_ctx *ctx, double *dev);
int b200_h2d(b200_ctx *ctx, double *dev, const double *host, long long n);
int b200_d2h(b200_ctx *ctx, double *host, const double *dev, long long n);
/* page-lock / unlock a caller-owned host array (cudaHostRegister) so host-vector entry
 * points copy at full PCIe speed; hypre allocates its vectors with plain malloc */
int b200_host_pin(void *host, long long bytes);
int b200_host_unpin(void *host);

/* ------------------------------------------------------------ BLAS-1 (dev)
 * hypre_SeqVectorSetConstantValues / Copy / Scale / Axpy / InnerProd
 * (src/seq_mv/vector.c:350-386,435-460,652-689,722-749,1069-1097) and
 * hypre_ParVectorInnerProd = local dot + Allreduce (src/parcsr_mv/par_vector.c:513-533). */
int b200_vec_fill(b200_ctx *ctx, int n, double a, double *y);
int b200_vec_copy(b200_ctx *ctx, int n, const double *x, double *y);
int b200_vec_scale(b200_ctx *ctx, int n, double a, double *y);
int b200_vec_axpy(b200_ctx *ctx, int n, double a, const double *x, double *y);
int b200_vec_dot(b200_ctx *ctx, int n, const double *x, const double *y, double *result_h);
/* NO COUNTERPART IN THE REFERENCE -- added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The local inner
 * product <x, y> of device vectors, written to the device word d_result on the context's stream; no host
 * synchronisation, no all-reduce, and NO allocation of any kind (no cudaMalloc / cudaFree, neither here nor in any
 * set-up this call needs): the block partials of the reduction live in a file-scope static device array of
 * B200_NOALLOC_DOT_MAX_BLOCKS doubles (128 KB, held by every GPU that loads the library until the process exits).
 * n > B200_NOALLOC_DOT_MAX_N is rejected with B200_ERR_ARG.  All calls share that array: calls that overlap on one
 * GPU (two streams, two contexts) race on it and must be serialised by the caller.  Deterministic for a given n;
 * the summation order differs from b200_vec_dot's (same tolerance against the exact sum). */
#define B200_NOALLOC_DOT_MAX_BLOCKS 16384
#define B200_NOALLOC_DOT_MAX_N      67108864      /* 16384 blocks x 4096 elements */
int b200_vec_dot_noalloc(b200_ctx *ctx, int n, const double *x, const double *y, double *d_result);

/* ------------------------------------------------------------------ ParCSR
 * Device mirror of hypre_ParCSRMatrix (src/parcsr_mv/par_csr_matrix.h:27-84):
 * diag (n_rows x n_cols_diag) and offd (n_rows x n_cols_offd, compressed
 * columns) CSR blocks, plus the halo plan of hypre_ParCSRCommPkg
 * (src/parcsr_mv/par_csr_communication.h:51-75).  All arrays are HOST arrays
 * and are copied; offd_* and the comm arrays may be NULL when n_cols_offd == 0
 * / num_sends == num_recvs == 0 (single rank). */
int b200_parcsr_create(b200_ctx *ctx,
                       int n_rows, int n_cols_diag,
                       const int *diag_i, const int *diag_j, const double *diag_a,
                       int n_cols_offd,
                       const int *offd_i, const int *offd_j, const double *offd_a,
                       int num_sends, const int *send_procs, const int *send_map_starts,
                       const int *send_map_elmts,
                       int num_recvs, const int *recv_procs, const int *recv_vec_starts,
                       b200_parcsr **A);
/* A level gathered on every rank (replicated coarse level of a multi-rank hierarchy): rows_per_block[b] = rows owned
 * by rank b.  Two-stage Gauss-Seidel then uses the strict lower triangle of the owners' diagonal blocks only, as the
 * reference's distributed sweep does (hypre_BoomerAMGRelaxTwoStageGaussSeidelHost, par_relax.c:1505-1588). */
int b200_parcsr_set_gs_blocks(b200_parcsr *A, int n_blocks, const int *rows_per_block);
int b200_parcsr_destroy(b200_parcsr *A);
int b200_parcsr_info(const b200_parcsr *A, int *n_rows, int *n_cols, long long *nnz,
                     int *layout /* 0 = CSR-vector, 1 = sliced-ELL, 2 = sliced-ELL with rows sorted in 256-row windows */, int *lanes_per_row);
/* Storage of the diag block.  Thread-per-row layouts (1, 2) are held compressed when that reads fewer
 * bytes (lossless; same products, same order as the CSR row of seq_mv/csr_matvec.c):
 *   value_format 4  pattern storage: one byte per entry = index of its (column - row, value) pair in a
 *                   dictionary of <= 254 pairs, and nothing per entry where the 32 rows of a slice share the
 *                   pattern (stencil operators and their first Galerkin products);
 *   value_format 1/2  16-bit column offsets from a base per (slice, entry) + 8 / 16-bit codes into a value
 *                   dictionary (<= 256 / 65536 distinct values);  3  16-bit columns + fp64 values;
 *   value_format 0  plain (int32 column + fp64 value).
 * stored_bytes = matrix bytes one pass reads; csr_bytes = 12 nnz + 4 (n+1), what the reference's CSR
 * (seq_mv/csr_matrix.h) would read. */
int b200_parcsr_storage(const b200_parcsr *A, int *value_format, long long *stored_bytes,
                        long long *csr_bytes, int *n_dictionary);
/* mode: -1 automatic (default; environment B200_COMPRESS=0 disables), 0 off, 1/2/3 force that value
 * format where the dictionary allows it, 4 the pattern storage where the block qualifies (tests/bench).
 * no_padding_markers (formats 1-3; set BEFORE the upload): 1 packs padding lanes as (own column) x
 * (dictionary entry 0.0) instead of a marker, which gives a predicate-free loop; 0 markers; -1 leaves it
 * (environment B200_ZPAD). */
int b200_set_compression(int mode, int no_padding_markers);
/* launch options of the thread-per-row kernels (A/B measurements): chunked 1 = every block walks one
 * contiguous run of 32-row slices (default; neighbouring rows meet in L1), 0 = grid stride, 2 = also the
 * value-dictionary / plain kernels; pat_regs: bits 1-3 = grid of the pattern kernel in waves of resident blocks
 * (8 = four waves, the default; 0 leaves it), bit 0 is ignored (it selected register-heavier instantiations of the
 * pattern kernel, which were measured slower on the B200 and removed); -1 leaves a value. */
int b200_set_kernel_options(int chunked, int pat_regs);
/* Test hook for the storage formats (host only, no device needed, never on a solve path): packs the CSR
 * block with the current layout / compression settings exactly as b200_parcsr_create would and computes
 * y = A x on the host FROM THE PACKED ARRAYS with the kernels' addressing (same fma chain per row). */
int b200_storage_decode_host(int n_rows, int n_cols, const int *i, const int *j, const double *a,
                             const double *x, double *y, int *layout, int *value_format,
                             long long *stored_bytes, long long *n_escape_groups);
/* force the storage layout chosen for the diag block (tests/bench): -1 auto */
int b200_set_layout_override(int layout, int lanes_per_row);

/* ------------------------------------------------------- row partitioning (host)
 * Splits a replicated GLOBAL CSR matrix into this rank's ParCSR blocks and halo plan --
 * what hypre_CSRMatrixToParCSRMatrix / GenerateDiagAndOffd
 * (src/parcsr_mv/par_csr_matrix.c:1812-2230) and hypre_MatvecCommPkgCreate
 * (src/parcsr_mv/par_csr_communication.c:1163-1200) produce, derived locally because every
 * rank sees the whole matrix (no MPI exists on the GPU box).  row_starts / col_starts have
 * nranks+1 entries.  Pure host code (works without a GPU). */
typedef struct b200_rowpart b200_rowpart;
int b200_rowpart_create(int rank, int nranks, int n_rows, int n_cols,
                        const int *gi, const int *gj, const double *ga,
                        const int *row_starts, const int *col_starts, b200_rowpart **part);
int b200_rowpart_destroy(b200_rowpart *part);
int b200_rowpart_sizes(const b200_rowpart *part, int *n_rows, int *n_cols_diag, int *n_cols_offd,
                       int *nnz_diag, int *nnz_offd, int *num_sends, int *n_send_elmts, int *num_recvs);
/* which: 0 diag_i, 1 diag_j, 2 offd_i, 3 offd_j, 4 col_map_offd, 5 send_procs,
 * 6 send_map_starts, 7 send_map_elmts, 8 recv_procs, 9 recv_vec_starts (borrowed pointers) */
int b200_rowpart_int_array(const b200_rowpart *part, int which, const int **data, int *len);
/* which: 0 diag values, 1 offd values */
int b200_rowpart_real_array(const b200_rowpart *part, int which, const double **data, int *len);
int b200_parcsr_create_from_rowpart(b200_ctx *ctx, const b200_rowpart *part, b200_parcsr **A);

/* y = alpha*A*x + beta*b   (dev pointers; y may alias b, must not alias x)
 * hypre_ParCSRMatrixMatvecOutOfPlace (src/parcsr_mv/par_csr_matvec.c:20-265)
 * over hypre_CSRMatrixMatvecOutOfPlace (src/seq_mv/csr_matvec.c:21-891). */
int b200_parcsr_matvec(b200_parcsr *A, double alpha, const double *x,
                       double beta, const double *b, double *y);
/* y = alpha*A^T*x + beta*y
 * hypre_ParCSRMatrixMatvecT (src/parcsr_mv/par_csr_matvec.c:287-546); the
 * explicit transpose is built once on first use (the reference transposes on
 * every call on its GPU path, src/seq_mv/csr_spmv_device.c:424-427). */
int b200_parcsr_matvecT(b200_parcsr *A, double alpha, const double *x,
                        double beta, double *y);
/* Same as b200_parcsr_matvec with HOST vectors: copies x (and b) to the device,
 * runs the kernel, copies y back.  This is the call the reference-facing
 * HYPRE_ParCSRMatrixMatvec binding makes for host-resident hypre vectors. */
int b200_parcsr_matvec_host(b200_parcsr *A, double alpha, const double *x_h,
                            double beta, const double *b_h, double *y_h);
/* The transposed product with HOST vectors (x_h: n_rows entries, y_h: this rank's n_cols_diag): the call the
 * HYPRE_ParCSRMatrixMatvecT binding makes (src/parcsr_mv/HYPRE_parcsr_matrix.c; hypre_ParCSRMatrixMatvecT,
 * src/parcsr_mv/par_csr_matvec.c:522-546). */
int b200_parcsr_matvecT_host(b200_parcsr *A, double alpha, const double *x_h, double beta, double *y_h);

/* --------------------------------------------------------------- smoothers
 * One sweep each, dev pointers.  `tmp` arguments are caller-provided work
 * vectors of length n_rows.  zero_guess != 0 promises u == 0 on entry (the
 * reference's hypre_ParVectorAllZeros flag, src/parcsr_mv/par_vector.h:24-44). */

/* relax_type 18 (and 7): u += w*(f - A u) ./ l1
 * hypre_BoomerAMGRelax7Jacobi (src/parcsr_ls/par_relax.c:1177-1254). */
int b200_relax_l1jacobi(b200_parcsr *A, const double *f, const double *l1, double w,
                        double *u, double *tmp, int zero_guess);
/* relax_type 16: hypre_ParCSRRelax_Cheby_SolveHost (src/parcsr_ls/par_cheby.c:224-385).
 * coefs_h = HOST array of order+1 coefficients... only the first `order` are used
 * exactly as the reference does (cheby_order = order-1). */
int b200_relax_cheby(b200_parcsr *A, const double *f, const double *ds, const double *coefs_h,
                     int order, int scale, double *u, double *v, double *r, double *tmp,
                     int zero_guess);
/* relax_type 11 (inner=1) / 12 (inner=2):
 * hypre_BoomerAMGRelaxTwoStageGaussSeidelHost (src/parcsr_ls/par_relax.c:1505-1588).
 * Returns B200_ERR_NUMERIC if a diagonal entry is zero (reference: hypre_error_in_arg(1)). */
int b200_relax_2stage_gs(b200_parcsr *A, const double *f, double w, int inner,
                         double *u, double *z, double *tmp, int zero_guess);

/* --------------------------------------------------------------- hierarchy
 * Device mirror of the solve-phase fields of hypre_ParAMGData
 * (src/parcsr_ls/par_amg.h:20-300).  The hierarchy is produced by the
 * reference's host setup and uploaded once. */
int b200_amg_create(b200_ctx *ctx, int num_levels, b200_amg **amg);
int b200_amg_destroy(b200_amg *amg);
/* A_array[level]; P_array[level] (NULL on the coarsest level).  The amg object
 * takes ownership of A and P. */
int b200_amg_set_level(b200_amg *amg, int level, b200_parcsr *A, b200_parcsr *P);
/* N > 1 ranks: levels >= `level` (>= 1) are replicated on every rank -- their A/P are uploaded as
 * whole single-rank matrices, P_array[level-1] with GLOBAL coarse columns -- so the small levels run
 * without any communication (reference: seq_threshold / hypre_seqAMGCycle,
 * src/parcsr_ls/par_cycle.c:362-369).  Call before b200_amg_finalize. */
int b200_amg_set_replicated_from(b200_amg *amg, int level);
/* grid_relax_type[1..3], num_grid_sweeps[1..3] (index = cycle_param: 1 down,
 * 2 up, 3 coarsest), cycle_type (1 = V, 2 = W), fcycle */
int b200_amg_set_cycle(b200_amg *amg, const int relax_type[4], const int num_sweeps[4],
                       int cycle_type, int fcycle);
/* per-level smoother data (HOST arrays, copied):
 * l1_norms[level] (relax 18/7: l1 row norms; 11/12: the diagonal), relax_weight[level] */
int b200_amg_set_level_relax(b200_amg *amg, int level, double relax_weight,
                             const double *l1_norms_h);
/* cheby_ds[level], cheby_coefs[level], cheby_order, cheby_scale
 * (src/parcsr_ls/par_cheby.c:57-206) */
int b200_amg_set_level_cheby(b200_amg *amg, int level, const double *ds_h,
                             const double *coefs_h, int order, int scale);
/* coarsest level, relax_type 9: the replicated dense N x N matrix A_mat (row
 * major) of hypre_GaussElimSetup (src/parcsr_ls/par_gauss_elim.c:32-454), this
 * rank's first row and the per-rank row counts for the gather. */
int b200_amg_set_coarse_dense(b200_amg *amg, int n_global, const double *A_mat_h,
                              int first_row, const int *rows_per_rank_h);
/* cycles per application when the hierarchy preconditions a Krylov solver (b200_pcg_solve, b200_gmres_solve,
 * b200_bicgstab_solve): hypre_ParAMGDataMaxIter with tol 0 -- hypre_BoomerAMGSolve then runs that many cycles
 * without residual norms (src/parcsr_ls/par_amg_solve.c:257-336; ij -precon_cycles).  Default 1. */
int b200_amg_set_precond_cycles(b200_amg *amg, int cycles);
/* CF-ordered relaxation (hypre_BoomerAMGRelaxIF, src/parcsr_ls/par_relax_interface.c:20-64; hypre_ParAMGDataRelaxOrder,
 * ij -CF 1): with relax_order 1 a sweep on a non-coarsest level is two calls of the smoother -- C points then F
 * points on the way down, F then C on the way up (relax types 0 / 7 / 18 through hypre_ParVectorElmdivpyMarked,
 * seq_mv/vector.c:823-870; two-stage Gauss-Seidel ignores the marker and runs twice; Chebyshev is not affected).
 * cf_marker_h: CF_marker_array[level] of the setup (this rank's rows; 1 = C point), required on every
 * non-coarsest level for the Jacobi types.  Both before b200_amg_finalize. */
int b200_amg_set_relax_order(b200_amg *amg, int relax_order);
int b200_amg_set_level_cf_marker(b200_amg *amg, int level, const int *cf_marker_h);
/* finalise: allocates F/U/work vectors, builds transposes and the launch plan */
int b200_amg_finalize(b200_amg *amg);
/* one multigrid cycle u <- u + B(f - A u)   (dev pointers, level-0 size)
 * hypre_BoomerAMGCycle (src/parcsr_ls/par_cycle.c:22-803).  zero_guess as above. */
int b200_amg_cycle(b200_amg *amg, const double *f, double *u, int zero_guess);
/* algorithmic HBM bytes of one cycle, from the uploaded level sizes */
int b200_amg_cycle_bytes(const b200_amg *amg, double *bytes);
/* the same model with the bytes the device layouts actually store (b200_parcsr_storage) */
int b200_amg_cycle_bytes_stored(const b200_amg *amg, double *bytes);
/* Persistent coarse-level kernel: the deepest run of levels whose operators are small CSR blocks without
 * communication (every level below ~4e6 nonzeros on one rank; the replicated levels on several) is executed as ONE
 * cooperative kernel per visit -- the operators of hypre_BoomerAMGCycle (par_cycle.c:209-779) on those levels in
 * order, a grid-wide barrier between them -- instead of one launch each.  Same arithmetic, bit-identical results.
 * b200_set_tail_limit: largest level (nonzeros of A_l) that joins, 0 = off (applies to hierarchies finalized
 * afterwards; environment B200_TAIL_NNZ); b200_amg_tail_level: first level inside the kernel, -1 = none.
 * Off by default: measured on the B200 it is slower than the graph-replayed launches. */
int b200_set_tail_limit(long long max_nnz);
int b200_amg_tail_level(const b200_amg *amg, int *level);
/* borrowed handle to A_array[level] / P_array[level] (owned by amg) */
int b200_amg_level_matrix(const b200_amg *amg, int level, b200_parcsr **A);
int b200_amg_level_interp(const b200_amg *amg, int level, b200_parcsr **P);
int b200_amg_level_info(const b200_amg *amg, int level, int *n_rows, long long *nnz_A,
                        long long *nnz_P);
/* relaxation sweeps per level of one cycle (host arithmetic, no device): what hypre_BoomerAMGCycle adds up
 * into the "cycle complexity" hypre_BoomerAMGSolve prints (src/parcsr_ls/par_cycle.c:417-430,
 * par_amg_solve.c:338-395).  num_sweeps[4] as in b200_amg_create; sweeps_per_level has num_levels entries. */
int b200_cycle_sweep_counts(int num_levels, int cycle_type, int fcycle, const int *num_sweeps,
                            long long *sweeps_per_level);
/* halo exchanges per level of one cycle (host arithmetic, no device): fwdA[l] forward exchanges (job 1 of
 * hypre_ParCSRCommHandleCreate, src/parcsr_mv/par_csr_communication.c:357-637) of the SpMV-type launches on A_l,
 * fwdP[l] forward exchanges on P_l (prolongation), revP[l] reverse exchanges (job 2) on P_l (restriction); arrays
 * of num_levels entries.  With b200_parcsr_halo_info they give the bytes a cycle puts on NVLink. */
int b200_cycle_exchange_counts(int num_levels, int cycle_type, int fcycle, const int *relax_type,
                               const int *num_sweeps, int cheby_order, int zero_guess,
                               long long *fwdA, long long *fwdP, long long *revP);
/* plan sizes of a block: messages / doubles per forward exchange (num_sends, n_send_elmts) and per reverse exchange
 * (num_recvs, n_cols_offd); any pointer may be NULL */
int b200_parcsr_halo_info(const b200_parcsr *A, int *num_sends, int *n_send_elmts, int *num_recvs,
                          int *n_cols_offd);
/* stand-alone AMG iteration (ij -solver 0): hypre_BoomerAMGSolve
 * (src/parcsr_ls/par_amg_solve.c:21-424) with converge_type 0. */
int b200_amg_solve(b200_amg *amg, const double *b, double *x, double tol, int min_iter,
                   int max_iter, int *num_iterations, double *rel_resid_norm,
                   double *resid_norms_h /* HOST, max_iter+1 entries ||r_k||_2, may be NULL */);

/* --------------------------------------------------------------------- PCG
 * hypre_PCGSolve (src/krylov/pcg.c:317-1008) with the ParCSR operator table
 * (src/parcsr_ls/par_krylov_func.c:14-330).  amg == NULL -> no preconditioner
 * (identity), amg != NULL -> one b200_amg_cycle per iteration
 * (HYPRE_BoomerAMGSolve with tol 0, max_iter 1).  b, x are dev pointers;
 * norms_h / rel_norms_h (HOST, length max_iter+1, may be NULL) receive the
 * history the reference logs (pcg.c:767-771). */
typedef struct {
   double tol;        /* r_tol  */
   double a_tol;      /* absolute tolerance */
   int    max_iter;
   int    two_norm;
   int    skip_break; /* pcg.c:581-629 */
   int    diag_scale; /* 1: preconditioner = diagonal scaling (HYPRE_ParCSRDiagScale) when amg == NULL */
} b200_pcg_params;

typedef struct {
   int    num_iterations;
   int    converged;
   int    error_flag;      /* HYPRE_ERROR_CONV (256) semantics: 0 or 256 */
   double rel_residual_norm;
   double bi_prod;
} b200_pcg_result;

int b200_pcg_solve(b200_ctx *ctx, b200_parcsr *A, b200_amg *amg, const b200_pcg_params *prm,
                   const double *b, double *x, b200_pcg_result *res,
                   double *norms_h, double *rel_norms_h);
/* same with HOST b/x: H2D of b and x, device solve, D2H of x (end-to-end call
 * made by the hypre_PCGSolve binding for host-resident hypre vectors) */
int b200_pcg_solve_host(b200_ctx *ctx, b200_parcsr *A, b200_amg *amg, const b200_pcg_params *prm,
                        const double *b_h, double *x_h, b200_pcg_result *res,
                        double *norms_h, double *rel_norms_h);

/* --------------------------------------------------- GMRES / COGMRES / BiCGSTAB
 * hypre_GMRESSolve (src/krylov/gmres.c:274-930; ij -solver 3 / 4), hypre_COGMRESSolve (src/krylov/cogmres.c;
 * ij -solver 16 / 17; b200_gmres_params.ortho = 1) and hypre_BiCGSTABSolve
 * (src/krylov/bicgstab.c:226-640; ij -solver 9 / 10) with the ParCSR operator table
 * (src/parcsr_ls/par_krylov_func.c:14-330): right preconditioning by one b200_amg_cycle (amg != NULL),
 * diagonal scaling (diag_scale, HYPRE_ParCSRDiagScale) or nothing.  Options the reference switches off by
 * default and that are not implemented must be off: rel_change, hybrid (GMRES: stop_crit).  b, x dev pointers;
 * norms_h (HOST, max_iter + 1 entries, may be NULL) receives the residual norms the reference logs
 * (gmres.c:413, 610; bicgstab.c:353, 489). */
typedef struct {
   double tol;               /* r_tol */
   double a_tol;
   int    max_iter;
   int    min_iter;
   int    k_dim;             /* restart length (HYPRE_GMRESSetKDim; ij default 5) */
   int    skip_real_r_check; /* gmres.c:767-771 */
   int    diag_scale;        /* as b200_pcg_params */
   int    ortho;             /* 0: modified Gram-Schmidt = hypre_GMRESSolve; 1: classical Gram-Schmidt with one
                              * multi-vector inner product and one multi-vector axpy per iteration =
                              * hypre_COGMRESSolve with cgs 1 (src/krylov/cogmres.c:289-…; MassInnerProd / MassAxpy,
                              * src/parcsr_ls/par_krylov_func.c:195-262; ij -solver 16 / 17) */
   int    flex;              /* 1: flexible GMRES = hypre_FlexGMRESSolve (src/krylov/flexgmres.c:282-…; ij -solver 61 / 60)
                              * with the default (empty) modify_pc: the preconditioned basis vectors are kept and
                              * the solution is updated from them without a further preconditioner application;
                              * true-residual check without the "residual no longer decreases" exit */
   double cf_tol;            /* > 0: abandon the solve when the weighted average convergence factor exceeds it
                              * (HYPRE_GMRESSetConvergenceFactorTol, gmres.c:632-652, 732-735); 0 = off */
} b200_gmres_params;

typedef struct {
   double tol;
   double a_tol;
   int    max_iter;
   int    min_iter;
   int    diag_scale;
   int    stop_crit;         /* 1: |r| <= tol (a_tol when set) as an absolute bound, bicgstab.c:372-383 */
   double cf_tol;            /* as b200_gmres_params (bicgstab.c:519-540) */
} b200_bicgstab_params;

/* hypre_LGMRESSolve (src/krylov/lgmres.c:337-...; ij -solver 51 / 50): GMRES(k_dim) augmented with the last
 * aug_dim error approximations (HYPRE_LGMRESSetAugDim, default 2); approx_constant = 1 keeps the space at k_dim
 * vectors (lgmres.c:105, the reference's fixed setting) */
typedef struct {
   double tol;
   double a_tol;
   int    max_iter;
   int    min_iter;
   int    k_dim;
   int    aug_dim;
   int    approx_constant;
   int    diag_scale;
   double cf_tol;            /* as b200_gmres_params (lgmres.c) */
} b200_lgmres_params;

typedef struct {
   int    num_iterations;
   int    converged;
   int    error_flag;        /* 0, or a sum of: 1 Inf/NaN in the input (HYPRE_ERROR_GENERIC), 2 BiCGSTAB
                              * breakdown (HYPRE_ERROR_GENERIC with message), 256 HYPRE_ERROR_CONV */
   double rel_residual_norm;
   double b_norm;
   int    real_residual_exit;/* 1: the solve ended on an evaluation of b - A x (the exits after which the
                              * reference prints "Final L2 norm of residual", gmres.c:512-520, 783-851;
                              * bicgstab.c:436-441, 508-512) */
} b200_krylov_result;

int b200_gmres_solve(b200_ctx *ctx, b200_parcsr *A, b200_amg *amg, const b200_gmres_params *prm,
                     const double *b, double *x, b200_krylov_result *res, double *norms_h);
int b200_bicgstab_solve(b200_ctx *ctx, b200_parcsr *A, b200_amg *amg, const b200_bicgstab_params *prm,
                        const double *b, double *x, b200_krylov_result *res, double *norms_h);
int b200_lgmres_solve(b200_ctx *ctx, b200_parcsr *A, b200_amg *amg, const b200_lgmres_params *prm,
                      const double *b, double *x, b200_krylov_result *res, double *norms_h);
/* the same with HOST b/x (H2D of b and x, device solve, D2H of x): the calls the hypre_GMRESSolve /
 * hypre_BiCGSTABSolve bindings make for host-resident hypre vectors */
int b200_gmres_solve_host(b200_ctx *ctx, b200_parcsr *A, b200_amg *amg, const b200_gmres_params *prm,
                          const double *b_h, double *x_h, b200_krylov_result *res, double *norms_h);
int b200_bicgstab_solve_host(b200_ctx *ctx, b200_parcsr *A, b200_amg *amg, const b200_bicgstab_params *prm,
                             const double *b_h, double *x_h, b200_krylov_result *res, double *norms_h);
int b200_lgmres_solve_host(b200_ctx *ctx, b200_parcsr *A, b200_amg *amg, const b200_lgmres_params *prm,
                           const double *b_h, double *x_h, b200_krylov_result *res, double *norms_h);

#ifdef __cplusplus
}
#endif
#endif /* B200AMG_H */
