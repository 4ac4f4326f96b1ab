// amg_solve.cu -- stand-alone AMG iteration (hypre_BoomerAMGSolve,
// src/parcsr_ls/par_amg_solve.c:165-345, converge_type 0): cycles until
// ||b - A x|| / ||b|| < tol.  The residual SpMV and its norm are one launch.
#include "ops.cuh"
#include "vecops.cuh"
#include "p2p.cuh"
#include <cmath>

// r = b - t ; red0 += r*r
struct EpiResidualNorm {
   static constexpr int NRED = 1;
   EPI_PLAIN_X
   const double *b; double *r;
   __device__ void operator()(int row, double t, double *red) const
   {
      double v = b[row] - t;
      r[row] = v;
      red[0] = fma(v, v, red[0]);
   }
};

enum { S_RES = 16, S_RHS = 17 };

static int residual_norm(b200_parcsr *A, const double *b, const double *x, double *r, double *nrm)
{
   b200_ctx *c = A->ctx;
   EpiResidualNorm e{b, r};
   B200_TRY(run_rows(A, A->diag, A->offd, true, x, e, c->d_scalars + S_RES));
   B200_TRY(scalars_allreduce(c, c->d_scalars + S_RES, 1));
   B200_TRY(scalars_to_host(c, S_RES, 1));
   *nrm = sqrt(c->h_scalars[S_RES]);
   return B200_OK;
}

extern "C" int b200_amg_solve(b200_amg *amg, const double *b, double *x, double tol, int min_iter,
                              int max_iter, int *num_iterations, double *rel_resid_norm,
                              double *resid_norms_h)
{
   ARG_CHECK(amg != nullptr && b != nullptr && x != nullptr && max_iter >= 0);
   b200_parcsr *A = nullptr;
   B200_TRY(b200_amg_level_matrix(amg, 0, &A));
   b200_ctx *c = A->ctx;
   const int n = A->n_rows;
   double *r = nullptr;
   B200_TRY(ctx_workspace(c, WS_AMG_R, n, &r));
   double resid = 1.0, rhs_norm = 0.0, rel = 1.0;
   int status = B200_OK, cycles = 0;
   const bool track = tol > 0.0 || resid_norms_h != nullptr;
   B200_TRY(ctx_rendezvous(c));
   if (track)
   {
      status = residual_norm(A, b, x, r, &resid);
      if (status == B200_OK) { status = vec_dot(c, n, b, b, c->d_scalars + S_RHS); }
      if (status == B200_OK) { status = scalars_allreduce(c, c->d_scalars + S_RHS, 1); }
      if (status == B200_OK) { status = scalars_to_host(c, S_RHS, 1); }
      if (status != B200_OK) { return status; }
      rhs_norm = sqrt(c->h_scalars[S_RHS]);
      if (resid != 0.0 && !(resid / resid == resid / resid))
      {
         b200_set_error("INFs and/or NaNs detected in input (par_amg_solve.c:193-216)");
         return B200_ERR_NUMERIC;
      }
      rel = rhs_norm != 0.0 ? resid / rhs_norm : resid;
      if (resid_norms_h) { resid_norms_h[0] = resid; }
   }
   while ((rel >= tol || cycles < min_iter) && cycles < max_iter)
   {
      status = b200_amg_cycle(amg, b, x, 0);
      if (status != B200_OK) { break; }
      if (track)
      {
         status = residual_norm(A, b, x, r, &resid);
         if (status != B200_OK) { break; }
         rel = rhs_norm != 0.0 ? resid / rhs_norm : resid;
         if (resid_norms_h) { resid_norms_h[cycles + 1] = resid; }
      }
      ++cycles;
   }
   if (num_iterations) { *num_iterations = cycles; }
   if (rel_resid_norm) { *rel_resid_norm = rel; }
   return status;
}
