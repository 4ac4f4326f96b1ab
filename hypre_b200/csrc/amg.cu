// amg.cu -- device multigrid cycle (hypre_BoomerAMGCycle, src/parcsr_ls/par_cycle.c:22-803).
#include "ops.cuh"
#include "vecops.cuh"
#include "p2p.cuh"
#include "tail.cuh"
#include <cmath>
#include <climits>

struct AmgLevel {
   b200_parcsr *A = nullptr, *P = nullptr;
   int     n = 0;
   double  relax_weight = 1.0;
   double *d_l1 = nullptr;
   double *d_ds = nullptr;
   int    *d_cf = nullptr;      // CF_marker_array[level] (relax_order 1), nullptr on the coarsest level
   std::vector<double> coefs;
   int     cheby_order = 2, cheby_scale = 1;
   double *F = nullptr, *U = nullptr, *Ualt = nullptr;   // F/U owned for level > 0, Ualt for all
};

struct b200_amg {
   b200_ctx *ctx = nullptr;
   int num_levels = 0;
   std::vector<AmgLevel> L;
   int relax_type[4] = {0, 18, 18, 9};
   int num_sweeps[4] = {1, 1, 1, 1};
   int cycle_type = 1, fcycle = 0;
   double *Vtemp = nullptr, *Ztemp = nullptr, *Ptemp = nullptr, *Rtemp = nullptr;
   // coarsest level dense solve (relax_type 9)
   int     coarse_n = 0, coarse_first_row = 0;
   std::vector<int> coarse_rows_per_rank;
   double *d_Ainv = nullptr;   // this rank's rows of A_mat^{-1}: [n_local x coarse_n]
   double *d_fall = nullptr;   // gathered coarse rhs [coarse_n]
   // levels >= repl_level are REPLICATED on every rank (full matrices, no communication): the
   // counterpart of the reference's seq_threshold / hypre_seqAMGCycle (src/parcsr_ls/par_cycle.c:362-369).
   // The level above holds P with global coarse columns; restriction onto repl_level is followed by
   // a vector all-reduce, prolongation from it is local.
   int  repl_level = 1 << 30;
   int  precond_cycles = 1;    // cycles per preconditioner application (hypre_ParAMGDataMaxIter with tol 0)
   int  relax_order = 0;       // hypre_ParAMGDataRelaxOrder: 1 = C then F points on the way down, F then C on the way up
   // levels >= tail_level run inside ONE persistent kernel per visit (tail.cuh): small CSR-vector levels without
   // communication (all of them on one rank, the replicated ones on several); INT_MAX = none
   int  tail_level = INT_MAX;
   bool finalized = false;
};

extern "C" int b200_amg_create(b200_ctx *ctx, int num_levels, b200_amg **out)
{
   ARG_CHECK(ctx != nullptr && out != nullptr && num_levels >= 1);
   b200_amg *a = new b200_amg();
   a->ctx = ctx; a->num_levels = num_levels; a->L.resize(num_levels);
   *out = a;
   return B200_OK;
}

extern "C" int b200_amg_destroy(b200_amg *a)
{
   if (!a) { return B200_OK; }
   cudaSetDevice(a->ctx->device);
   cudaStreamSynchronize(a->ctx->stream);
   tail_drop_programs(a->ctx);       // recorded programs point into this hierarchy
   a->ctx->generation++;
   for (auto &l : a->L)
   {
      b200_parcsr_destroy(l.A); b200_parcsr_destroy(l.P);
      cudaFree(l.d_l1); cudaFree(l.d_ds); cudaFree(l.d_cf); cudaFree(l.F); cudaFree(l.U); cudaFree(l.Ualt);
   }
   cudaFree(a->Vtemp); cudaFree(a->Ztemp); cudaFree(a->Ptemp); cudaFree(a->Rtemp);
   cudaFree(a->d_Ainv); cudaFree(a->d_fall);
   delete a;
   return B200_OK;
}

extern "C" int b200_amg_set_level(b200_amg *a, int level, b200_parcsr *A, b200_parcsr *P)
{
   ARG_CHECK(a != nullptr && level >= 0 && level < a->num_levels && A != nullptr);
   ARG_CHECK((level == a->num_levels - 1) == (P == nullptr));
   a->L[level].A = A; a->L[level].P = P; a->L[level].n = A->n_rows;
   return B200_OK;
}

extern "C" int b200_amg_set_replicated_from(b200_amg *a, int level)
{
   ARG_CHECK(a != nullptr && level >= 1);
   a->repl_level = level;
   return B200_OK;
}

extern "C" int b200_amg_set_cycle(b200_amg *a, const int relax_type[4], const int num_sweeps[4],
                                  int cycle_type, int fcycle)
{
   ARG_CHECK(a != nullptr && relax_type != nullptr && num_sweeps != nullptr && cycle_type >= 1);
   for (int k = 0; k < 4; k++) { a->relax_type[k] = relax_type[k]; a->num_sweeps[k] = num_sweeps[k]; }
   a->cycle_type = cycle_type; a->fcycle = fcycle;
   a->ctx->generation++;
   return B200_OK;
}

extern "C" int b200_amg_set_level_relax(b200_amg *a, int level, double relax_weight, const double *l1_h)
{
   ARG_CHECK(a != nullptr && level >= 0 && level < a->num_levels && a->L[level].A != nullptr);
   AmgLevel &l = a->L[level];
   a->ctx->generation++;
   l.relax_weight = relax_weight;
   cudaFree(l.d_l1); l.d_l1 = nullptr;
   if (l1_h) { B200_TRY(dev_upload(&l.d_l1, l1_h, (size_t)l.n)); }
   return B200_OK;
}

extern "C" int b200_amg_set_level_cheby(b200_amg *a, int level, const double *ds_h, const double *coefs_h,
                                        int order, int scale)
{
   ARG_CHECK(a != nullptr && level >= 0 && level < a->num_levels && a->L[level].A != nullptr);
   ARG_CHECK(coefs_h != nullptr && order >= 1);
   AmgLevel &l = a->L[level];
   a->ctx->generation++;
   cudaFree(l.d_ds); l.d_ds = nullptr;
   if (ds_h) { B200_TRY(dev_upload(&l.d_ds, ds_h, (size_t)l.n)); }
   ARG_CHECK(!scale || ds_h != nullptr);
   l.coefs.assign(coefs_h, coefs_h + (order > 4 ? 4 : order) + 1);
   l.cheby_order = order; l.cheby_scale = scale;
   return B200_OK;
}

// CF-ordered relaxation (hypre_BoomerAMGRelaxIF, src/parcsr_ls/par_relax_interface.c:20-64; ij -CF 1): with
// relax_order 1 every sweep on a non-coarsest level is two calls of the smoother, on the C points then the F points
// on the way down and F then C on the way up (Jacobi types 0 / 7 / 18: hypre_ParVectorElmdivpyMarked on CF_marker;
// the two-stage Gauss-Seidel sweeps ignore the marker and simply run twice; Chebyshev does not go through RelaxIF).
extern "C" int b200_amg_set_relax_order(b200_amg *a, int relax_order)
{
   ARG_CHECK(a != nullptr && (relax_order == 0 || relax_order == 1));
   a->relax_order = relax_order;
   a->ctx->generation++;
   return B200_OK;
}

extern "C" int b200_amg_set_level_cf_marker(b200_amg *a, int level, const int *cf_marker_h)
{
   ARG_CHECK(a != nullptr && level >= 0 && level < a->num_levels && a->L[level].A != nullptr);
   AmgLevel &l = a->L[level];
   a->ctx->generation++;
   cudaFree(l.d_cf); l.d_cf = nullptr;
   if (cf_marker_h) { B200_TRY(dev_upload(&l.d_cf, cf_marker_h, (size_t)l.n)); }
   return B200_OK;
}

// A_mat^{-1} by the reference's own elimination (hypre_gselim, src/utilities/gselim.h:11-67:
// no pivoting, forward elimination then back substitution) applied to the N unit vectors.
static int invert_like_gselim(int n, const double *A_in, std::vector<double> *inv)
{
   std::vector<double> A(A_in, A_in + (size_t)n * n);
   std::vector<double> X((size_t)n * n, 0.0);   // X[row][rhs]
   for (int i = 0; i < n; i++) { X[(size_t)i * n + i] = 1.0; }
   if (n == 1)
   {
      if (A[0] == 0.0) { return 1; }
      X[0] = 1.0 / A[0];
      *inv = X;
      return 0;
   }
   for (int k = 0; k < n - 1; k++)
   {
      if (A[(size_t)k * n + k] != 0.0)
      {
         double divA = 1.0 / A[(size_t)k * n + k];
         for (int j = k + 1; j < n; j++)
         {
            if (A[(size_t)j * n + k] != 0.0)
            {
               double factor = A[(size_t)j * n + k] * divA;
               for (int m = k + 1; m < n; m++) { A[(size_t)j * n + m] -= factor * A[(size_t)k * n + m]; }
               for (int c = 0; c < n; c++) { X[(size_t)j * n + c] -= factor * X[(size_t)k * n + c]; }
            }
         }
      }
   }
   for (int k = n - 1; k > 0; --k)
   {
      if (A[(size_t)k * n + k] != 0.0)
      {
         for (int c = 0; c < n; c++) { X[(size_t)k * n + c] /= A[(size_t)k * n + k]; }
         for (int j = 0; j < k; j++)
         {
            if (A[(size_t)j * n + k] != 0.0)
            {
               for (int c = 0; c < n; c++) { X[(size_t)j * n + c] -= X[(size_t)k * n + c] * A[(size_t)j * n + k]; }
            }
         }
      }
   }
   if (A[0] != 0.0) { for (int c = 0; c < n; c++) { X[c] /= A[0]; } }
   *inv = X;
   return 0;
}

extern "C" int b200_amg_set_coarse_dense(b200_amg *a, int n_global, const double *A_mat_h, int first_row,
                                         const int *rows_per_rank_h)
{
   ARG_CHECK(a != nullptr && n_global >= 0);
   b200_ctx *c = a->ctx;
   a->coarse_n = n_global; a->coarse_first_row = first_row;
   a->coarse_rows_per_rank.assign(c->nranks, 0);
   if (rows_per_rank_h) { a->coarse_rows_per_rank.assign(rows_per_rank_h, rows_per_rank_h + c->nranks); }
   else if (c->nranks == 1) { a->coarse_rows_per_rank[0] = n_global; }
   else { b200_set_error("rows_per_rank required for nranks > 1"); return B200_ERR_ARG; }
   cudaFree(a->d_Ainv); cudaFree(a->d_fall); a->d_Ainv = nullptr; a->d_fall = nullptr;
   if (n_global == 0) { return B200_OK; }
   const int n_local = a->coarse_rows_per_rank[c->rank];
   // a rank without coarse rows holds no matrix (the reference gathers it on the owners only,
   // par_gauss_elim.c:60-200); it still takes part in the sum of the right-hand side
   ARG_CHECK(n_local == 0 || A_mat_h != nullptr);
   if (n_local > 0)
   {
      std::vector<double> inv;
      if (invert_like_gselim(n_global, A_mat_h, &inv) != 0)
      {
         b200_set_error("coarse matrix is singular (hypre_gselim error)");
         return B200_ERR_NUMERIC;
      }
      B200_TRY(dev_upload(&a->d_Ainv, inv.data() + (size_t)first_row * n_global, (size_t)n_local * n_global));
   }
   CUDA_TRY(cudaMalloc((void **)&a->d_fall, sizeof(double) * (size_t)n_global));
   return B200_OK;
}

// largest level (in nonzeros of A) that joins the persistent coarse-level kernel; 0 switches the kernel off
static long long g_tail_nnz = []() { const char *e = getenv("B200_TAIL_NNZ"); return e ? atoll(e) : 0LL; }();
extern "C" int b200_set_tail_limit(long long max_nnz)
{
   ARG_CHECK(max_nnz >= 0);
   g_tail_nnz = max_nnz;
   return B200_OK;
}

static bool relax_supported(int t) { return t == 18 || t == 7 || t == 16 || t == 11 || t == 12 || t == 9; }

extern "C" int b200_amg_finalize(b200_amg *a)
{
   ARG_CHECK(a != nullptr);
   b200_ctx *c = a->ctx;
   int maxn = 1;
   for (int l = 0; l < a->num_levels; l++)
   {
      AmgLevel &L = a->L[l];
      ARG_CHECK(L.A != nullptr);
      maxn = L.n > maxn ? L.n : maxn;
      if (l < a->num_levels - 1)
      {
         ARG_CHECK(L.P != nullptr && L.P->n_rows == L.n);
         ARG_CHECK(L.P->n_cols_diag == a->L[l + 1].n);
         B200_TRY(parcsr_ensure_T(L.P));
      }
   }
   for (int k = 1; k <= 3; k++)
   {
      if (!relax_supported(a->relax_type[k]))
      {
         b200_set_error("relax_type %d (cycle_param %d) has no device implementation", a->relax_type[k], k);
         return B200_ERR_UNSUPPORTED;
      }
   }
   if (a->relax_type[1] == 9 || a->relax_type[2] == 9)
   {
      b200_set_error("Gaussian elimination is only supported on the coarsest level");
      return B200_ERR_UNSUPPORTED;
   }
   if (a->relax_type[3] == 9 && a->d_Ainv == nullptr && a->L[a->num_levels - 1].n > 0)
   {
      b200_set_error("coarsest level uses relax_type 9 but b200_amg_set_coarse_dense was not called");
      return B200_ERR_ARG;
   }
   for (int l = 0; l < a->num_levels; l++)
   {
      AmgLevel &L = a->L[l];
      const int cp = (l == a->num_levels - 1 && a->num_levels > 1) ? 3 : 1;
      for (int k = (cp == 3 ? 3 : 1); k <= (cp == 3 ? 3 : 2); k++)
      {
         const int t = a->relax_type[k];
         if ((t == 18 || t == 7) && L.d_l1 == nullptr) { b200_set_error("level %d: l1 norms missing", l); return B200_ERR_ARG; }
         if (t == 16 && L.coefs.empty()) { b200_set_error("level %d: Chebyshev data missing", l); return B200_ERR_ARG; }
         if (t == 11 || t == 12) { B200_TRY(parcsr_ensure_L(L.A)); }
         if (a->relax_order == 1 && cp != 3 && a->num_levels > 1 && (t == 18 || t == 7) && L.d_cf == nullptr && L.n > 0)
         {
            b200_set_error("level %d: relax_order 1 needs the CF marker (b200_amg_set_level_cf_marker)", l);
            return B200_ERR_ARG;
         }
      }
      if (l > 0)
      {
         B200_TRY(b200_dev_alloc(c, L.n, &L.F));
         B200_TRY(b200_dev_alloc(c, L.n, &L.U));
      }
      B200_TRY(b200_dev_alloc(c, L.n, &L.Ualt));
   }
   B200_TRY(b200_dev_alloc(c, maxn, &a->Vtemp));
   B200_TRY(b200_dev_alloc(c, maxn, &a->Ztemp));
   B200_TRY(b200_dev_alloc(c, maxn, &a->Ptemp));
   B200_TRY(b200_dev_alloc(c, maxn, &a->Rtemp));
   // persistent coarse-level kernel: the deepest run of levels whose operators are CSR-vector blocks without
   // communication and small enough that launch latency, not bandwidth, is what they cost
   {
      const long long tail_nnz = g_tail_nnz;
      const int nl = a->num_levels;
      a->tail_level = INT_MAX;
      bool coarse_ok = c->nranks == 1 || a->repl_level <= nl - 1;       // no gather of the coarse right-hand side
      auto quiet = [](const b200_parcsr *M) { return M->num_sends + M->num_recvs == 0; };
      for (int l = nl - 1; l >= 1 && coarse_ok && tail_nnz > 0; l--)
      {
         const AmgLevel &L = a->L[l];
         bool ok = L.A->diag.layout == 0 && quiet(L.A) && L.A->diag.nnz <= tail_nnz && (!L.A->hasL || L.A->lower.layout == 0);
         if (l < nl - 1) { ok = ok && L.P->diag.layout == 0 && L.P->diagT.layout == 0 && quiet(L.P); }
         if (!ok) { break; }
         a->tail_level = l;
      }
      // a tail of one level (the dense solve alone) gains nothing
      if (a->tail_level >= nl - 1) { a->tail_level = INT_MAX; }
   }
   a->finalized = true;
   return B200_OK;
}

// ---------------------------------------------------------------- the cycle
struct CycleState {
   std::vector<double *> buf[2];   // the two U buffers of each level
   std::vector<int>      cur;      // which one currently holds U[l]
   std::vector<char>     zero;     // hypre_ParVectorAllZeros(U[l])
};

static inline double *U_cur(CycleState &s, int l)   { return s.buf[s.cur[l]][l]; }
static inline double *U_other(CycleState &s, int l) { return s.buf[1 - s.cur[l]][l]; }

static int materialize_zero(b200_amg *a, CycleState &s, int l)
{
   if (s.zero[l]) { B200_TRY(vec_fill(a->ctx, a->L[l].n, 0.0, U_cur(s, l))); s.zero[l] = 0; }
   return B200_OK;
}

static int coarse_dense_solve(b200_amg *a, const double *f, double *u)
{
   b200_ctx *c = a->ctx;
   const int n_local = a->coarse_rows_per_rank[c->rank];
   if (a->coarse_n == 0) { return B200_OK; }
   if (c->nranks > 1 && a->repl_level > a->num_levels - 1)
   {
      // hypre_MPI_Allgatherv of the coarse rhs (src/parcsr_ls/par_gauss_elim.c:575-576):
      // every rank contributes its slice into a zeroed vector, summed over ranks
      CUDA_TRY(cudaMemsetAsync(a->d_fall, 0, sizeof(double) * (size_t)a->coarse_n, c->stream));
      if (n_local > 0)
      {
         CUDA_TRY(cudaMemcpyAsync(a->d_fall + a->coarse_first_row, f, sizeof(double) * (size_t)n_local,
                                  cudaMemcpyDeviceToDevice, c->stream));
      }
      if (c->use_p2p && a->coarse_n <= 256) { B200_TRY(p2p_allreduce(c, a->d_fall, a->coarse_n)); }
      else { NCCL_TRY(ncclAllReduce(a->d_fall, a->d_fall, (size_t)a->coarse_n, ncclDouble, ncclSum, c->comm, c->stream)); }
      return op_dense_matvec(c, n_local, a->coarse_n, a->d_Ainv, a->d_fall, u);
   }
   return op_dense_matvec(c, n_local, a->coarse_n, a->d_Ainv, f, u);
}

static int relax_once(b200_amg *a, CycleState &s, int l, int relax_type, const double *f, const FusedDot *fd = nullptr,
                      int points = 0)
{
   AmgLevel &L = a->L[l];
   const int zg = s.zero[l] ? 1 : 0;
   switch (relax_type)
   {
      case 18: case 7:
         if (zg) { B200_TRY(op_l1jacobi(L.A, f, L.d_l1, L.relax_weight, nullptr, U_cur(s, l), 1, nullptr, L.d_cf, points)); }
         else
         {
            B200_TRY(op_l1jacobi(L.A, f, L.d_l1, L.relax_weight, U_cur(s, l), U_other(s, l), 0, fd, L.d_cf, points));
            s.cur[l] = 1 - s.cur[l];
         }
         break;
      case 16:
         // hypre_ParCSRRelax_Cheby_Solve(A, f, ds, coefs, order, scale, variant, u, Vtemp, Ztemp, Ptemp, Rtemp)
         // (par_cycle.c:529-538)
         B200_TRY(op_cheby(L.A, f, L.d_ds, L.coefs.data(), L.cheby_order, L.cheby_scale, U_cur(s, l),
                           a->Vtemp, a->Ztemp, a->Rtemp, zg, fd));
         break;
      case 11: case 12:
         B200_TRY(op_gs2(L.A, f, L.relax_weight, relax_type == 11 ? 1 : 2, U_cur(s, l), a->Vtemp, a->Ztemp, zg, fd));
         break;
      case 9:
         B200_TRY(coarse_dense_solve(a, f, U_cur(s, l)));
         break;
      default:
         b200_set_error("relax_type %d unsupported", relax_type);
         return B200_ERR_UNSUPPORTED;
   }
   s.zero[l] = 0;   // hypre_BoomerAMGRelax clears all_zeros (par_relax.c:170)
   return B200_OK;
}

// One cycle; the level-counter state machine is the reference's (par_cycle.c:209-223,647-779).
// u_alt0: second level-0 buffer.  On return *u_result points at the buffer holding U[0].
// fd (optional): <fd->with, U[0]> is reduced inside the cycle's last smoothing launch on level 0 where the
// smoother supports it (*fd->done tells); the caller (PCG: gamma = <r, C r>) falls back to its own pass otherwise.
static int amg_cycle_body(b200_amg *a, const double *f, double *u, int zero_guess, double **u_result,
                          const FusedDot *fd);

static int amg_cycle_core(b200_amg *a, const double *f, double *u, int zero_guess, double **u_result,
                          const FusedDot *fd = nullptr)
{
   const int st = amg_cycle_body(a, f, u, zero_guess, u_result, fd);
   if (st != B200_OK && tail_recording(a->ctx)) { tail_drop_programs(a->ctx); }     // leave the recorder clean
   return st;
}

static int amg_cycle_body(b200_amg *a, const double *f, double *u, int zero_guess, double **u_result,
                          const FusedDot *fd)
{
   const int nl = a->num_levels;
   CycleState s;
   s.buf[0].resize(nl); s.buf[1].resize(nl); s.cur.assign(nl, 0); s.zero.assign(nl, 0);
   std::vector<const double *> F(nl);
   for (int l = 0; l < nl; l++)
   {
      s.buf[0][l] = l == 0 ? u : a->L[l].U;
      s.buf[1][l] = a->L[l].Ualt;
      F[l] = l == 0 ? f : a->L[l].F;
   }
   s.zero[0] = zero_guess ? 1 : 0;
   if (zero_guess && nl > 1)
   {
      // out-of-place Jacobi sweeps swap the two level-0 buffers; with a zero initial guess the
      // first write may go to either buffer, so start where an odd number of swaps ends in `u`
      auto jac = [](int t) { return t == 18 || t == 7; };
      int swaps = 0; bool z = true;
      const int calls = a->relax_order == 1 ? 2 : 1;       // CF ordering: two calls per sweep
      for (int j = 0; j < a->num_sweeps[1] * calls; j++) { if (jac(a->relax_type[1]) && !z) { swaps++; } z = false; }
      for (int j = 0; j < a->num_sweeps[2] * calls; j++) { if (jac(a->relax_type[2])) { swaps++; } }
      if (swaps & 1) { s.cur[0] = 1; }
   }

   std::vector<int> lev_counter(nl);
   lev_counter[0] = 1;
   for (int k = 1; k < nl; k++) { lev_counter[k] = a->fcycle ? 1 : a->cycle_type; }
   int fcycle_lev = nl - 2;
   int level = 0, cycle_param = 1;
   bool not_finished = true;

   const int Lt = a->tail_level;
   while (not_finished)
   {
      // levels >= Lt: operators are recorded, the whole visit runs as one persistent kernel (tail.cuh)
      if (level >= Lt && !tail_recording(a->ctx)) { B200_TRY(tail_begin(a->ctx)); }
      int relax_type, num_sweep;
      if (nl > 1) { relax_type = a->relax_type[cycle_param]; num_sweep = a->num_sweeps[cycle_param]; }
      else        { relax_type = a->relax_type[0];           num_sweep = a->num_sweeps[0]; }
      // level 0 is visited twice (lev_counter 1 -> 0 -> -1): its last sweep of the second visit ends the cycle
      const bool final_visit = level == 0 && (nl == 1 || lev_counter[0] == 0);
      // hypre_BoomerAMGRelaxIF: relax_order 1 splits a sweep into two calls (C, F down; F, C up)
      const bool cf_order = a->relax_order == 1 && cycle_param < 3 && nl > 1 && relax_type != 16 && relax_type != 9;
      for (int j = 0; j < num_sweep; j++)
      {
         for (int pass = 0; pass < (cf_order ? 2 : 1); pass++)
         {
            const int points = !cf_order ? 0 : (((cycle_param < 2) == (pass == 0)) ? 1 : -1);
            const bool last_call = final_visit && j == num_sweep - 1 && pass == (cf_order ? 1 : 0);
            B200_TRY(relax_once(a, s, level, relax_type, F[level], last_call ? fd : nullptr, points));
         }
      }

      --lev_counter[level];
      if (lev_counter[level] >= 0 && level != nl - 1)
      {
         // residual + restriction (par_cycle.c:662-705)
         const int fine = level, coarse = level + 1;
         AmgLevel &L = a->L[fine];
         s.zero[coarse] = 1;   // hypre_ParVectorSetZeros(U_array[coarse_grid])
         s.cur[coarse] = 0;
         const double *r = a->Vtemp;
         if (s.zero[fine]) { r = F[fine]; }   // f - A*0 == f
         else { B200_TRY(op_matvec(L.A, -1.0, U_cur(s, fine), 1.0, F[fine], a->Vtemp)); }
         B200_TRY(op_matvecT(L.P, 1.0, r, 0.0, a->L[coarse].F));
         if (coarse == a->repl_level && a->ctx->nranks > 1)
         {
            // every rank restricted its own rows onto the whole coarse level: sum over ranks
            b200_ctx *c = a->ctx;
            const int nc = a->L[coarse].n;
            if (c->use_p2p && nc <= p2p_allreduce_vec_max()) { B200_TRY(p2p_allreduce_vec(c, a->L[coarse].F, nc)); }
            else { NCCL_TRY(ncclAllReduce(a->L[coarse].F, a->L[coarse].F, (size_t)nc, ncclDouble, ncclSum, c->comm, c->stream)); }
         }
         ++level;
         lev_counter[level] = lev_counter[level] > a->cycle_type ? lev_counter[level] : a->cycle_type;
         cycle_param = (level == nl - 1) ? 3 : 1;
      }
      else if (level != 0)
      {
         // interpolation (par_cycle.c:751-757): U[fine] += P U[coarse]
         const int fine = level - 1, coarse = level;
         AmgLevel &L = a->L[fine];
         if (fine < Lt && tail_recording(a->ctx)) { B200_TRY(tail_flush(a->ctx)); }      // leaving the tail
         B200_TRY(materialize_zero(a, s, coarse));
         if (s.zero[fine])
         {
            B200_TRY(op_matvec(L.P, 1.0, U_cur(s, coarse), 0.0, nullptr, U_cur(s, fine)));
            s.zero[fine] = 0;
         }
         else
         {
            B200_TRY(op_matvec(L.P, 1.0, U_cur(s, coarse), 1.0, U_cur(s, fine), U_cur(s, fine)));
         }
         --level;
         cycle_param = 2;
         if (a->fcycle && fcycle_lev == level)
         {
            lev_counter[level] = lev_counter[level] > 1 ? lev_counter[level] : 1;
            fcycle_lev--;
         }
      }
      else { not_finished = false; }
   }
   if (tail_recording(a->ctx)) { B200_TRY(tail_flush(a->ctx)); }
   B200_TRY(materialize_zero(a, s, 0));
   *u_result = U_cur(s, 0);
   return B200_OK;
}

int amg_cycle_ptr(b200_amg *a, const double *f, double *u, int zero_guess, double **u_result)
{
   return amg_cycle_core(a, f, u, zero_guess, u_result);
}
double *amg_level0_alt(b200_amg *a) { return a->L[0].Ualt; }

// BoomerAMG as a preconditioner: z = B r.  hypre_BoomerAMGSolve with tol 0 and max_iter k (ij -precon_cycles k)
// runs k cycles without looking at the residual (src/parcsr_ls/par_amg_solve.c:257-336); the Krylov solver cleared
// u before the call, so the first cycle starts from a zero guess and the others continue on its result.
int amg_precond_ptr(b200_amg *a, const double *f, double *u, double **u_result, const FusedDot *fd)
{
   B200_TRY(amg_cycle_core(a, f, u, 1, u_result, a->precond_cycles == 1 ? fd : nullptr));
   for (int k = 1; k < a->precond_cycles; k++)
   {
      if (*u_result != u) { B200_TRY(b200_vec_copy(a->ctx, a->L[0].n, *u_result, u)); }
      B200_TRY(amg_cycle_core(a, f, u, 0, u_result, k == a->precond_cycles - 1 ? fd : nullptr));
   }
   return B200_OK;
}
int amg_precond_cycles(const b200_amg *a) { return a->precond_cycles; }

extern "C" int b200_amg_set_precond_cycles(b200_amg *a, int cycles)
{
   ARG_CHECK(a != nullptr && cycles >= 1);
   a->precond_cycles = cycles;
   a->ctx->generation++;
   return B200_OK;
}

extern "C" int b200_amg_cycle(b200_amg *a, const double *f, double *u, int zero_guess)
{
   ARG_CHECK(a != nullptr && a->finalized && f != nullptr && u != nullptr);
   double *res = nullptr;
   B200_TRY(amg_cycle_core(a, f, u, zero_guess, &res));
   if (res != u) { B200_TRY(b200_vec_copy(a->ctx, a->L[0].n, res, u)); }
   return B200_OK;
}

// first level of the persistent coarse-level kernel (-1: none)
extern "C" int b200_amg_tail_level(const b200_amg *a, int *level)
{
   ARG_CHECK(a != nullptr && level != nullptr);
   *level = a->tail_level == INT_MAX ? -1 : a->tail_level;
   return B200_OK;
}

extern "C" int b200_amg_level_matrix(const b200_amg *a, int level, b200_parcsr **A)
{
   ARG_CHECK(a != nullptr && A != nullptr && level >= 0 && level < a->num_levels);
   *A = a->L[level].A;
   return B200_OK;
}
extern "C" int b200_amg_level_interp(const b200_amg *a, int level, b200_parcsr **P)
{
   ARG_CHECK(a != nullptr && P != nullptr && level >= 0 && level < a->num_levels);
   *P = a->L[level].P;
   return B200_OK;
}

extern "C" int b200_amg_level_info(const b200_amg *a, int level, int *n_rows, long long *nnz_A, long long *nnz_P)
{
   ARG_CHECK(a != nullptr && level >= 0 && level < a->num_levels);
   const AmgLevel &L = a->L[level];
   if (n_rows) { *n_rows = L.n; }
   if (nnz_A) { *nnz_A = L.A ? L.A->diag.nnz + L.A->offd.nnz : 0; }
   if (nnz_P) { *nnz_P = L.P ? L.P->diag.nnz + L.P->offd.nnz : 0; }
   return B200_OK;
}

// Algorithmic bytes of one cycle (SURVEY.md section 8(d) formulas, int32 + fp64):
//   SpMV-type pass over A: 12 nnz + 4 n + vectors (8 bytes per vector element streamed)
static inline double dmat_pass_bytes(const DMat &D, bool stored)
{
   if (stored && D.layout >= 1) { return (double)D.stored_bytes; }
   return 12.0 * (double)D.nnz + 4.0 * (double)D.n_rows;
}

// stored = false: the reference's CSR bytes (the algorithmic figure); stored = true: the bytes the
// device layout of each block (compressed sliced ELL where chosen) makes one pass read
static int cycle_bytes(const b200_amg *a, bool stored, double *bytes)
{
   ARG_CHECK(a != nullptr && bytes != nullptr);
   double tot = 0.0;
   const int nl = a->num_levels;
   for (int l = 0; l < nl; l++)
   {
      const AmgLevel &L = a->L[l];
      const double n = L.n, nnz = (double)(L.A->diag.nnz + L.A->offd.nnz);
      const double spmv = dmat_pass_bytes(L.A->diag, stored) + 12.0 * (double)L.A->offd.nnz;
      auto sweep_bytes = [&](int t, bool zero) -> double {
         switch (t)
         {
            case 18: case 7: return zero ? 24.0 * n : spmv + 32.0 * n;       // u,f,l1 in, u out
            case 16: {
               int k = (L.cheby_order > 4 ? 4 : L.cheby_order) - 1;
               if (zero && k >= 1)
               {
                  // fused first step (EpiChebyZero): f, ds in, u (or out, r) out; then k-1 plain steps
                  return spmv + 24.0 * n + (k > 1 ? 16.0 * n : 0.0) + (k - 1) * (spmv + 40.0 * n);
               }
               double first = zero ? 32.0 * n : spmv + 40.0 * n;            // u,f,ds in; r,tmp out
               return first + k * (spmv + 40.0 * n);                         // tmp,r,ds,u in; u/out
            }
            case 11: case 12: {
               int inner = t == 11 ? 1 : 2;
               double low = L.A->hasL ? dmat_pass_bytes(L.A->lower, stored) : 6.0 * nnz + 4.0 * n;
               double first = zero ? 24.0 * n : spmv + 32.0 * n;            // u,f,d in; z out
               return first + inner * (low + 40.0 * n);                      // z,d,u in; z,u out
            }
            default: return 0.0;
         }
      };
      if (l < nl - 1)
      {
         const double nc = a->L[l + 1].n;
         const double passP  = dmat_pass_bytes(L.P->diag, stored) + 12.0 * (double)L.P->offd.nnz;
         const double passPT = (L.P->hasT ? dmat_pass_bytes(L.P->diagT, stored)
                                          : 12.0 * (double)L.P->diag.nnz + 4.0 * nc) + 12.0 * (double)L.P->offd.nnz;
         // V-cycle: pre sweeps (first one with zero guess), residual, restriction, prolongation, post sweeps
         const int calls = (a->relax_order == 1 && a->relax_type[1] != 16) ? 2 : 1;
         const int calls2 = (a->relax_order == 1 && a->relax_type[2] != 16) ? 2 : 1;
         for (int j = 0; j < a->num_sweeps[1] * calls; j++) { tot += sweep_bytes(a->relax_type[1], j == 0); }
         tot += spmv + 24.0 * n;                          // residual: u,f in; r out
         tot += passPT + 8.0 * n + 8.0 * nc;                    // restriction
         tot += passP + 8.0 * nc + 16.0 * n;                    // prolongation
         for (int j = 0; j < a->num_sweeps[2] * calls2; j++) { tot += sweep_bytes(a->relax_type[2], false); }
      }
      else if (a->relax_type[3] == 9) { tot += 8.0 * (double)a->coarse_n * a->coarse_n; }
      else { for (int j = 0; j < a->num_sweeps[3]; j++) { tot += sweep_bytes(a->relax_type[3], j == 0); } }
   }
   *bytes = tot;
   return B200_OK;
}

extern "C" int b200_amg_cycle_bytes(const b200_amg *a, double *bytes) { return cycle_bytes(a, false, bytes); }
extern "C" int b200_amg_cycle_bytes_stored(const b200_amg *a, double *bytes) { return cycle_bytes(a, true, bytes); }

// Relaxation sweeps per level in ONE cycle: a dry run of the level-counter state machine of amg_cycle_core
// (hypre_BoomerAMGCycle, par_cycle.c:200-300, 640-780) that launches nothing.  The binding needs it for the
// "cycle complexity" line hypre_BoomerAMGSolve prints (par_amg_solve.c:338-395: sum over the sweeps of the
// nonzeros of their level / nonzeros of level 0; par_cycle.c:417-430).  Pure host arithmetic.
extern "C" int b200_cycle_sweep_counts(int num_levels, int cycle_type, int fcycle, const int *num_sweeps,
                                       long long *sweeps_per_level)
{
   ARG_CHECK(num_levels >= 1 && cycle_type >= 1 && num_sweeps != nullptr && sweeps_per_level != nullptr);
   const int nl = num_levels;
   for (int l = 0; l < nl; l++) { sweeps_per_level[l] = 0; }
   std::vector<int> lev_counter(nl);
   lev_counter[0] = 1;
   for (int k = 1; k < nl; k++) { lev_counter[k] = fcycle ? 1 : cycle_type; }
   int fcycle_lev = nl - 2, level = 0, cycle_param = 1;
   for (;;)
   {
      sweeps_per_level[level] += nl > 1 ? num_sweeps[cycle_param] : num_sweeps[0];
      --lev_counter[level];
      if (lev_counter[level] >= 0 && level != nl - 1)
      {
         ++level;
         lev_counter[level] = lev_counter[level] > cycle_type ? lev_counter[level] : cycle_type;
         cycle_param = (level == nl - 1) ? 3 : 1;
      }
      else if (level != 0)
      {
         --level;
         cycle_param = 2;
         if (fcycle && fcycle_lev == level)
         {
            lev_counter[level] = lev_counter[level] > 1 ? lev_counter[level] : 1;
            fcycle_lev--;
         }
      }
      else { break; }
   }
   return B200_OK;
}

// Halo exchanges per level in ONE cycle, from the cycle parameters alone: the same dry run of the level-counter
// state machine as above, tracking the all-zeros flag of U as amg_cycle_core does.  fwdA[l] = forward exchanges
// (ghost values of the operand, hypre_ParCSRCommHandleCreate job 1) of SpMV-type launches on A_l (smoother sweeps,
// residual); fwdP[l] = forward exchanges on P_l (prolongation from level l+1); revP[l] = reverse exchanges
// (job 2) on P_l (restriction onto level l+1).  bench.py turns them into bytes on the wire with the plan sizes of
// b200_parcsr_halo_info: the NVLink side of the roofline.  Pure host arithmetic.
static long long sweep_exchanges(int t, bool zero, int cheby_order)
{
   switch (t)
   {
      case 18: case 7: return zero ? 0 : 1;                       // zero guess: u = w f ./ l1, no SpMV
      case 16: { const int k = (cheby_order > 4 ? 4 : cheby_order < 1 ? 1 : cheby_order) - 1; return zero ? k : k + 1; }
      case 11: case 12: return zero ? 0 : 1;                      // the strict-lower passes are rank-local
      default: return 0;                                          // 9: dense solve (gather of the coarse rhs)
   }
}

extern "C" int b200_cycle_exchange_counts(int num_levels, int cycle_type, int fcycle, const int *relax_type,
                                          const int *num_sweeps, int cheby_order, int zero_guess,
                                          long long *fwdA, long long *fwdP, long long *revP)
{
   ARG_CHECK(num_levels >= 1 && cycle_type >= 1 && relax_type != nullptr && num_sweeps != nullptr);
   ARG_CHECK(fwdA != nullptr && fwdP != nullptr && revP != nullptr);
   const int nl = num_levels;
   for (int l = 0; l < nl; l++) { fwdA[l] = fwdP[l] = revP[l] = 0; }
   std::vector<int> lev_counter(nl), zero(nl, 0);
   lev_counter[0] = 1;
   for (int k = 1; k < nl; k++) { lev_counter[k] = fcycle ? 1 : cycle_type; }
   zero[0] = zero_guess ? 1 : 0;
   int fcycle_lev = nl - 2, level = 0, cycle_param = 1;
   for (;;)
   {
      const int t = nl > 1 ? relax_type[cycle_param] : relax_type[0];
      const int ns = nl > 1 ? num_sweeps[cycle_param] : num_sweeps[0];
      for (int j = 0; j < ns; j++) { fwdA[level] += sweep_exchanges(t, zero[level] != 0, cheby_order); zero[level] = 0; }
      --lev_counter[level];
      if (lev_counter[level] >= 0 && level != nl - 1)
      {
         if (!zero[level]) { fwdA[level] += 1; }      // residual r = f - A u (skipped while u == 0: r = f)
         revP[level] += 1;                            // restriction
         ++level;
         zero[level] = 1;
         lev_counter[level] = lev_counter[level] > cycle_type ? lev_counter[level] : cycle_type;
         cycle_param = (level == nl - 1) ? 3 : 1;
      }
      else if (level != 0)
      {
         fwdP[level - 1] += 1;                        // prolongation
         zero[level - 1] = 0;
         --level;
         cycle_param = 2;
         if (fcycle && fcycle_lev == level)
         {
            lev_counter[level] = lev_counter[level] > 1 ? lev_counter[level] : 1;
            fcycle_lev--;
         }
      }
      else { break; }
   }
   return B200_OK;
}

// sizes of the halo plan of a ParCSR block (hypre_ParCSRCommPkg): messages and doubles per forward exchange
// (num_sends, n_send_elmts) and per reverse exchange (num_recvs, n_cols_offd)
extern "C" int b200_parcsr_halo_info(const b200_parcsr *A, int *num_sends, int *n_send_elmts, int *num_recvs,
                                     int *n_cols_offd)
{
   ARG_CHECK(A != nullptr);
   if (num_sends) { *num_sends = A->num_sends; }
   if (n_send_elmts) { *n_send_elmts = A->n_send_elmts; }
   if (num_recvs) { *num_recvs = A->num_recvs; }
   if (n_cols_offd) { *n_cols_offd = A->n_cols_offd; }
   return B200_OK;
}
