// pcg.cu -- preconditioned conjugate gradients on the device
// (hypre_PCGSolve, src/krylov/pcg.c:317-1008, with the ParCSR operator table of
// src/parcsr_ls/par_krylov_func.c).  Scalars stay on the device between the
// kernels of one iteration; the host reads {sdotp, <r,r>, <r,Cr>, breakdown flag}
// once per iteration for the reference's convergence / breakdown logic.
#include "ops.cuh"
#include "vecops.cuh"
#include "p2p.cuh"
#include <cmath>
#include <cfloat>

int     amg_cycle_ptr(b200_amg *a, const double *f, double *u, int zero_guess, double **u_result);

// scalar slots in ctx->d_scalars
enum { S_SDOTP = 8, S_IPROD = 9, S_GAMMA = 10, S_FLAG = 11, S_TMP = 12 };

// x += alpha p ; r -= alpha s ; red0 = <r,r>   with alpha = gamma / sdotp computed from the
// device-resident sdotp; the reference's breakdown guards (pcg.c:581-629) are evaluated
// here as well so that x and r stay untouched when the reference would `break`.
struct FPcgUpdate {
   static constexpr int NRED = 1;
   double gamma; int skip_break; const double *scal; const double *p; const double *s; double *x; double *r;
   __device__ void operator()(int i, double *red) const
   {
      const double sdotp = scal[S_SDOTP];
      const double alpha = gamma / sdotp;
      bool stop = (sdotp == 0.0);
      if (!stop)
      {
         if (alpha <= 0.0)               { stop = skip_break < 3; }
         else if (!(alpha >= 4.9406564584124654e-324)) { stop = skip_break < 2; }
         else if (!(alpha >= DBL_MIN))   { stop = skip_break < 1; }
      }
      double ri = r[i];
      if (!stop)
      {
         x[i] = fma(alpha, p[i], x[i]);
         ri = fma(-alpha, s[i], ri);
         r[i] = ri;
      }
      red[0] = fma(ri, ri, red[0]);
   }
};

// p = z + beta p   (pcg.c:960-971: ScaleVector(beta,p); Axpy(1,s,p))
struct FPcgDirection {
   static constexpr int NRED = 0;
   double beta; const double *z; double *p;
   __device__ void operator()(int i, double *) const { p[i] = beta * p[i] + z[i]; }
};

// z = r ./ d  (HYPRE_ParCSRDiagScale) and red0 = <r,z>
struct FDiagScaleDot {
   static constexpr int NRED = 1;
   const double *d; const double *r; double *z;
   __device__ void operator()(int i, double *red) const
   {
      double zi = r[i] / d[i];
      z[i] = zi;
      red[0] = fma(r[i], zi, red[0]);
   }
};

static int dot_to(b200_ctx *c, int n, const double *x, const double *y, int slot)
{
   return vec_dot(c, n, x, y, c->d_scalars + slot);
}

// z = C r.  Returns in *z_out the buffer that holds the result (z or the AMG's level-0 alternate).
static int apply_precond(b200_ctx *c, b200_parcsr *A, b200_amg *amg, const b200_pcg_params *prm,
                         const double *r, double *z, double **z_out)
{
   const int n = A->n_rows;
   *z_out = z;
   if (amg) { return amg_cycle_ptr(amg, r, z, 1, z_out); }
   if (prm->diag_scale)
   {
      B200_TRY(parcsr_ensure_L(A));
      FDiagScaleDot f{A->d_diagonal, r, z};
      return run_ewise(c, n, f, c->d_scalars + S_TMP);
   }
   return b200_vec_copy(c, n, r, z);   // identity: hypre_ParKrylovIdentity
}

extern "C" int b200_pcg_solve(b200_ctx *c, b200_parcsr *A, b200_amg *amg, const b200_pcg_params *prm,
                              const double *b, double *x, b200_pcg_result *res,
                              double *norms_h, double *rel_norms_h)
{
   ARG_CHECK(c != nullptr && A != nullptr && prm != nullptr && b != nullptr && x != nullptr && res != nullptr);
   ARG_CHECK(prm->max_iter >= 0);
   const int n = A->n_rows;
   const int two_norm = prm->two_norm;
   double *h = c->h_scalars;
   memset(res, 0, sizeof(*res));

   double *p = nullptr, *s = nullptr, *r = nullptr;
   B200_TRY(ctx_workspace(c, WS_PCG_P, n, &p));
   B200_TRY(ctx_workspace(c, WS_PCG_S, n, &s));
   B200_TRY(ctx_workspace(c, WS_PCG_R, n, &r));
   int status = B200_OK;
   int i = 0;
   double bi_prod = 0.0, i_prod = 0.0, i_prod_0 = 0.0, gamma = 0.0, gamma_old = 0.0, eps = 0.0;
   double *z = nullptr;

#define PCG_TRY(expr) do { status = (expr); if (status != B200_OK) goto done; } while (0)

   PCG_TRY(ctx_rendezvous(c));
   // bi_prod = <b,b>  (two_norm)  |  <C b, b>
   if (two_norm) { PCG_TRY(dot_to(c, n, b, b, S_TMP)); }
   else
   {
      PCG_TRY(apply_precond(c, A, amg, prm, b, p, &z));
      PCG_TRY(dot_to(c, n, z, b, S_TMP));
   }
   PCG_TRY(scalars_allreduce(c, c->d_scalars + S_TMP, 1));
   PCG_TRY(scalars_to_host(c, S_TMP, 1));
   bi_prod = h[S_TMP];
   res->bi_prod = bi_prod;
   if (bi_prod != 0.0 && !(bi_prod / bi_prod == bi_prod / bi_prod))
   {
      // INFs and/or NaNs detected in input (pcg.c:425-445)
      res->error_flag = 1;
      goto done;
   }
   eps = prm->tol * prm->tol;
   if (bi_prod > 0.0)
   {
      double e2 = prm->a_tol * prm->a_tol / bi_prod;
      eps = eps > e2 ? eps : e2;
   }
   else
   {
      // zero right-hand side: x = b (pcg.c:472-485)
      PCG_TRY(b200_vec_copy(c, n, b, x));
      if (norms_h) { norms_h[0] = 0.0; }
      if (rel_norms_h) { rel_norms_h[0] = 0.0; }
      goto done;
   }

   // r = b - A x ; p = C r ; gamma = <r,p>
   PCG_TRY(op_matvec(A, -1.0, x, 1.0, b, r));
   PCG_TRY(apply_precond(c, A, amg, prm, r, p, &z));
   if (z != p) { PCG_TRY(b200_vec_copy(c, n, z, p)); }
   PCG_TRY(dot_to(c, n, r, p, S_GAMMA));
   PCG_TRY(dot_to(c, n, r, r, S_IPROD));
   PCG_TRY(scalars_allreduce(c, c->d_scalars + S_IPROD, 2));
   PCG_TRY(scalars_to_host(c, S_IPROD, 2));
   gamma = h[S_GAMMA];
   if (gamma != 0.0 && !(gamma / gamma == gamma / gamma)) { res->error_flag = 1; goto done; }
   i_prod_0 = two_norm ? h[S_IPROD] : gamma;
   if (norms_h) { norms_h[0] = sqrt(i_prod_0); }

   while ((i + 1) <= prm->max_iter)
   {
      i++;
      // s = A p ; sdotp = <s,p>
      PCG_TRY(op_matvec_dot(A, p, s, c->d_scalars + S_SDOTP));
      PCG_TRY(scalars_allreduce(c, c->d_scalars + S_SDOTP, 1));
      // x += alpha p ; r -= alpha s ; <r,r>
      {
         FPcgUpdate f{gamma, prm->skip_break, c->d_scalars, p, s, x, r};
         PCG_TRY(run_ewise(c, n, f, c->d_scalars + S_IPROD));
      }
      // s = C r ; gamma = <r,s>
      PCG_TRY(apply_precond(c, A, amg, prm, r, s, &z));
      if (!amg && prm->diag_scale)
      {
         CUDA_TRY(cudaMemcpyAsync(c->d_scalars + S_GAMMA, c->d_scalars + S_TMP, sizeof(double),
                                  cudaMemcpyDeviceToDevice, c->stream));
      }
      else { PCG_TRY(dot_to(c, n, r, z, S_GAMMA)); }
      PCG_TRY(scalars_allreduce(c, c->d_scalars + S_IPROD, 2));
      PCG_TRY(scalars_to_host(c, S_SDOTP, 3));

      // ---- host-side logic of the reference on the three scalars
      {
         const double sdotp = h[S_SDOTP];
         bool brk = false;
         if (sdotp == 0.0) { res->error_flag |= 256; brk = true; }
         else
         {
            const double alpha = gamma / sdotp;
            if (alpha <= 0.0)                                  { res->error_flag |= 256; brk = prm->skip_break < 3; }
            else if (!(alpha >= 4.9406564584124654e-324))      { res->error_flag |= 256; brk = prm->skip_break < 2; }
            else if (!(alpha >= DBL_MIN))                      { res->error_flag |= 256; brk = prm->skip_break < 1; }
         }
         if (brk) { if (i == 1) { i_prod = i_prod_0; } break; }
      }
      gamma_old = gamma;
      gamma = h[S_GAMMA];
      i_prod = two_norm ? h[S_IPROD] : gamma;
      if (norms_h) { norms_h[i] = sqrt(i_prod); }
      if (rel_norms_h) { rel_norms_h[i] = bi_prod ? sqrt(i_prod / bi_prod) : 0.0; }

      if (i_prod / bi_prod < eps) { res->converged = 1; break; }

      if (gamma <= 0.0)                                   { res->error_flag |= 256; if (prm->skip_break < 3) { if (i == 1) { i_prod = i_prod_0; } break; } }
      else if (!(gamma >= 4.9406564584124654e-324))       { res->error_flag |= 256; if (prm->skip_break < 2) { if (i == 1) { i_prod = i_prod_0; } break; } }
      else if (!(gamma >= DBL_MIN))                       { res->error_flag |= 256; if (prm->skip_break < 1) { if (i == 1) { i_prod = i_prod_0; } break; } }

      // p = s + beta p
      {
         FPcgDirection f{gamma / gamma_old, z, p};
         PCG_TRY(run_ewise(c, n, f, nullptr));
      }
   }
   if (i >= prm->max_iter && (i_prod / bi_prod) >= eps && eps > 0) { res->error_flag |= 256; }
   res->num_iterations = i;
   res->rel_residual_norm = bi_prod > 0.0 ? sqrt(i_prod / bi_prod) : 0.0;

done:
#undef PCG_TRY
   cudaStreamSynchronize(c->stream);
   return status;
}

extern "C" int b200_pcg_solve_host(b200_ctx *c, b200_parcsr *A, b200_amg *amg, const b200_pcg_params *prm,
                                   const double *b_h, double *x_h, b200_pcg_result *res,
                                   double *norms_h, double *rel_norms_h)
{
   ARG_CHECK(c != nullptr && A != nullptr && b_h != nullptr && x_h != nullptr);
   const int n = A->n_rows;
   double *b = nullptr, *x = nullptr;
   B200_TRY(ctx_workspace(c, WS_HOST_B, n, &b));
   B200_TRY(ctx_workspace(c, WS_HOST_X, n, &x));
   CUDA_TRY(cudaMemcpyAsync(b, b_h, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
   CUDA_TRY(cudaMemcpyAsync(x, x_h, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
   B200_TRY(b200_pcg_solve(c, A, amg, prm, b, x, res, norms_h, rel_norms_h));
   return b200_d2h(c, x_h, x, n);
}
