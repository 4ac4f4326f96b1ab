// krylov_core.hpp -- the scalar logic of the restarted-GMRES family and of BiCGSTAB over a vector backend.
//
//   gmres_core     hypre_GMRESSolve      src/krylov/gmres.c:274-930       (ortho = 0, flex = 0)
//                  hypre_COGMRESSolve    src/krylov/cogmres.c, cgs == 1   (ortho = 1)
//                  hypre_FlexGMRESSolve  src/krylov/flexgmres.c           (flex = 1, default modify_pc)
//   lgmres_core    hypre_LGMRESSolve     src/krylov/lgmres.c
//   bicgstab_core  hypre_BiCGSTABSolve   src/krylov/bicgstab.c:226-640
//
// In the reference these are, like hypre_PCGSolve, host loops over an operator table (hypre_GMRESFunctions,
// gmres.h:64-90).  Here the loops are templates over a backend `B` whose methods are the (fused) vector
// operations; libb200amg instantiates them with the device backend of krylov.cu only.  The header has no CUDA
// in it, so the CPU tests can instantiate the very same loops over a host backend
// (tests/hostbackend/krylov_host_backend.cpp, test infrastructure) and compare every iterate with the oracle
// without a GPU.
//
// Backend concept (all vectors are `double *` handles of length ops.n owned by the caller; every method
// returns B200_OK or a negative error code which the loops pass straight up):
//   matvec(alpha, x, beta, b, y)        y = alpha A x + beta b            (b may be nullptr when beta == 0)
//   precond(r, z, &zr)                  zr -> C r ; zr is z or a buffer owned by the preconditioner that
//                                       stays valid until the next precond call (or r itself for identity)
//   copy(x, y)  scale(a, y)  axpy(a, x, y)      hypre_SeqVectorCopy / Scale / Axpy (src/seq_mv/vector.c)
//   dots(k, xs, ys, out)                out[q] = <xs[q], ys[q]> over all ranks, q < k <= 4, values on the HOST
//   mgs(i, p, h)                        modified Gram-Schmidt of p[i] against p[0..i-1] in that order:
//                                       h[j] = <p[j], p[i]>, p[i] -= h[j] p[j]  (j = 0..i-1), then
//                                       h[i] = <p[i], p[i]>; h on the HOST        (gmres.c:552-563)
//   cgs(i, p, h)                        classical Gram-Schmidt (COGMRES, cogmres.c:553-577 with cgs == 1): h[j] =
//                                       <p[i], p[j]> for all j < i from the SAME p[i] (MassInnerProd), then
//                                       p[i] += sum_j (-h[j]) p[j] in the order j = 0..i-1 (MassAxpy), then
//                                       h[i] = <p[i], p[i]>
//   bicg_half(alpha, v, q, x, r)        x += alpha v ; r -= alpha q                (bicgstab.c:439-440)
//   bicg_full(gamma, v, s, r0, x, r, out)   x += gamma v ; r -= gamma s ; out = {<r,r>, <r0,r>}   (:461-464, :525)
//   bicg_dir(gamma, f, q, r, p)         p = f (p - gamma q) + r                    (:528-540)
#pragma once
#include <cmath>
#include <cfloat>
#include <cstring>
#include <vector>
#include "../../include/b200amg.h"

#define KRY_TRY(expr) do { int _s = (expr); if (_s != B200_OK) return _s; } while (0)

// the reference's "adequate progress" test (gmres.c:632-652, bicgstab.c:519-540): true when the weighted average
// convergence factor exceeds cf_tol
static inline bool kry_cf_exceeded(double cf_tol, double r_norm, double r_norm_0, int iter, double *cf_ave_0, double *cf_ave_1)
{
   *cf_ave_0 = *cf_ave_1;
   *cf_ave_1 = pow(r_norm / r_norm_0, 1.0 / (2.0 * iter));
   double weight = fabs(*cf_ave_1 - *cf_ave_0);
   weight = weight / (*cf_ave_1 > *cf_ave_0 ? *cf_ave_1 : *cf_ave_0);
   weight = 1.0 - weight;
   return weight * *cf_ave_1 > cf_tol;
}

// true when v is an Inf or a NaN (the reference's ieee_check: v / v != v / v, gmres.c:378-404)
static inline bool kry_not_finite(double v) { return v != 0.0 && !(v / v == v / v); }

// Right-preconditioned restarted GMRES(k_dim) with rel_change = 0 and cf_tol = 0; prm->ortho == 1 gives
// hypre_COGMRESSolve with cgs == 1, which differs from hypre_GMRESSolve in the orthogonalisation only
// (and stores the same Hessenberg matrix column-major).
// p: k_dim + 1 basis vectors; r, w: work vectors.  norms_h (HOST, max_iter + 1 entries) may be nullptr.
// prm->flex == 1 gives hypre_FlexGMRESSolve (flexgmres.c): pp = k_dim further vectors that keep C p[j].
template <class B>
int gmres_core(B &ops, const b200_gmres_params *prm, const double *b, double *x, double *const *p,
               double *r, double *w, b200_krylov_result *res, double *norms_h, double *const *pp = nullptr)
{
   const bool flex = prm->flex != 0;
   const int k_dim = prm->k_dim, max_iter = prm->max_iter, min_iter = prm->min_iter;
   const double epsmac = 1.e-16;
   std::vector<double> rs(k_dim + 1, 0.0), c(k_dim, 0.0), s(k_dim, 0.0), col(k_dim + 2, 0.0);
   std::vector<double> hh((size_t)(k_dim + 1) * k_dim, 0.0);
#define HH(a, bb) hh[(size_t)(a) * k_dim + (bb)]
   double d2[2] = {0.0, 0.0};
   double b_norm, r_norm, den_norm, epsilon, t, gamma, real_r_norm_old, real_r_norm_new;
   double r_norm_0, cf_ave_0 = 0.0, cf_ave_1 = 0.0;
   bool break_value = false;
   int i = 0, j, k, iter = 0;
   const double *z = nullptr;
   memset(res, 0, sizeof(*res));

   // p[0] = b - A x ; ||b||, ||p[0]||   (gmres.c:364-415)
   KRY_TRY(ops.matvec(-1.0, x, 1.0, b, p[0]));
   {
      const double *xs[2] = {b, p[0]}, *ys[2] = {b, p[0]};
      KRY_TRY(ops.dots(2, xs, ys, d2));
   }
   b_norm = sqrt(d2[0]);
   real_r_norm_old = b_norm;
   res->b_norm = b_norm;
   if (kry_not_finite(b_norm)) { res->error_flag = 1; return B200_OK; }
   r_norm = sqrt(d2[1]);
   if (kry_not_finite(r_norm)) { res->error_flag = 1; return B200_OK; }
   r_norm_0 = r_norm;
   if (norms_h) { norms_h[0] = r_norm; }
   den_norm = b_norm > 0.0 ? b_norm : r_norm;
   epsilon = prm->a_tol > prm->tol * den_norm ? prm->a_tol : prm->tol * den_norm;

   while (iter < max_iter)
   {
      rs[0] = r_norm;
      if (r_norm == 0.0) { res->num_iterations = iter; return B200_OK; }      // gmres.c:484-497
      if (r_norm <= epsilon && iter >= min_iter)                              // :501-529
      {
         KRY_TRY(ops.matvec(-1.0, x, 1.0, b, r));
         const double *xs[1] = {r};
         KRY_TRY(ops.dots(1, xs, xs, d2));
         r_norm = sqrt(d2[0]);
         if (r_norm <= epsilon) { res->real_residual_exit = 1; break; }
      }
      t = 1.0 / r_norm;
      KRY_TRY(ops.scale(t, p[0]));
      i = 0;
      while (i < k_dim && iter < max_iter)                                    // restart cycle, :536-726
      {
         i++; iter++;
         if (flex)
         {
            // flexgmres.c:470-480: the preconditioned vector is stored (pre_vecs[i-1])
            KRY_TRY(ops.precond(p[i - 1], pp[i - 1], &z));
            if (z != pp[i - 1]) { KRY_TRY(ops.copy(z, pp[i - 1])); }
            z = pp[i - 1];
         }
         else { KRY_TRY(ops.precond(p[i - 1], r, &z)); }
         KRY_TRY(ops.matvec(1.0, z, 0.0, nullptr, p[i]));
         if (prm->ortho == 1) { KRY_TRY(ops.cgs(i, p, col.data())); }
         else                 { KRY_TRY(ops.mgs(i, p, col.data())); }
         for (j = 0; j < i; j++) { HH(j, i - 1) = col[j]; }
         t = sqrt(col[i]);
         HH(i, i - 1) = t;
         if (t != 0.0) { t = 1.0 / t; KRY_TRY(ops.scale(t, p[i])); }
         for (j = 1; j < i; j++)
         {
            t = HH(j - 1, i - 1);
            HH(j - 1, i - 1) = s[j - 1] * HH(j, i - 1) + c[j - 1] * t;
            HH(j, i - 1) = -s[j - 1] * t + c[j - 1] * HH(j, i - 1);
         }
         t = HH(i, i - 1) * HH(i, i - 1);
         t += HH(i - 1, i - 1) * HH(i - 1, i - 1);
         gamma = sqrt(t);
         if (gamma == 0.0) { gamma = epsmac; }
         c[i - 1] = HH(i - 1, i - 1) / gamma;
         s[i - 1] = HH(i, i - 1) / gamma;
         rs[i] = -HH(i, i - 1) * rs[i - 1];
         rs[i] /= gamma;
         rs[i - 1] = c[i - 1] * rs[i - 1];
         HH(i - 1, i - 1) = s[i - 1] * HH(i, i - 1) + c[i - 1] * HH(i - 1, i - 1);
         r_norm = fabs(rs[i]);
         if (norms_h) { norms_h[iter] = r_norm; }
         if (prm->cf_tol > 0.0 && kry_cf_exceeded(prm->cf_tol, r_norm, r_norm_0, iter, &cf_ave_0, &cf_ave_1)) { break_value = true; break; }
         if (r_norm <= epsilon && iter >= min_iter) { break; }
      }

      if (break_value) { break; }      // cf_tol: leave without the update of this cycle (:732-735)
      // upper triangular solve, x += C (sum_j rs[j] p[j])   (:735-762)
      rs[i - 1] = rs[i - 1] / HH(i - 1, i - 1);
      for (k = i - 2; k >= 0; k--)
      {
         t = 0.0;
         for (j = k + 1; j < i; j++) { t -= HH(k, j) * rs[j]; }
         t += rs[k];
         rs[k] = t / HH(k, k);
      }
      if (flex)
      {
         // linear combination of the stored preconditioned vectors (flexgmres.c:575-590)
         KRY_TRY(ops.copy(pp[i - 1], w));
         KRY_TRY(ops.scale(rs[i - 1], w));
         for (j = i - 2; j >= 0; j--) { KRY_TRY(ops.axpy(rs[j], pp[j], w)); }
         KRY_TRY(ops.axpy(1.0, w, x));
      }
      else
      {
         KRY_TRY(ops.copy(p[i - 1], w));
         KRY_TRY(ops.scale(rs[i - 1], w));
         for (j = i - 2; j >= 0; j--) { KRY_TRY(ops.axpy(rs[j], p[j], w)); }
         KRY_TRY(ops.precond(w, r, &z));
         KRY_TRY(ops.axpy(1.0, z, x));
      }

      // convergence: evaluate the actual residual (:765-871)
      if (r_norm <= epsilon && iter >= min_iter)
      {
         if (prm->skip_real_r_check && !flex) { res->converged = 1; break; }
         KRY_TRY(ops.matvec(-1.0, x, 1.0, b, r));
         const double *xs[1] = {r};
         KRY_TRY(ops.dots(1, xs, xs, d2));
         real_r_norm_new = r_norm = sqrt(d2[0]);
         if (r_norm <= epsilon) { res->converged = 1; res->real_residual_exit = 1; break; }
         // the true residual disagrees: stop if it no longer decreases, else restart from it
         if (!flex && real_r_norm_new >= real_r_norm_old) { res->converged = 1; res->real_residual_exit = 1; break; }
         KRY_TRY(ops.copy(r, p[0]));
         i = 0;
         real_r_norm_old = real_r_norm_new;
      }

      // residual vector of the restart (:874-890)
      for (j = i; j > 0; j--)
      {
         rs[j - 1] = -s[j - 1] * rs[j];
         rs[j] = c[j - 1] * rs[j];
      }
      if (i) { KRY_TRY(ops.axpy(rs[i] - 1.0, p[i], p[i])); }
      for (j = i - 1; j > 0; j--) { KRY_TRY(ops.axpy(rs[j], p[j], p[i])); }
      if (i)
      {
         KRY_TRY(ops.axpy(rs[0] - 1.0, p[0], p[0]));
         KRY_TRY(ops.axpy(1.0, p[i], p[0]));
      }
   }
#undef HH
   res->num_iterations = iter;
   res->rel_residual_norm = b_norm > 0.0 ? r_norm / b_norm : r_norm;
   if (iter >= max_iter && r_norm > epsilon && epsilon > 0) { res->error_flag |= 256; }   // HYPRE_ERROR_CONV, :908-911
   return B200_OK;
}

// BiCGSTAB with stop_crit = 0 and cf_tol = 0.  r, r0, s, v, p, q: work vectors.
// res->error_flag: 1 = Inf/NaN input, 2 = breakdown ("BiCGSTAB broke down", hypre_error_w_msg GENERIC), 256 = CONV.
template <class B>
int bicgstab_core(B &ops, const b200_bicgstab_params *prm, const double *b, double *x, double *r, double *r0,
                  double *s, double *v, double *p, double *q, b200_krylov_result *res, double *norms_h)
{
   const int max_iter = prm->max_iter, min_iter = prm->min_iter;
   const double epsmac = DBL_MIN;    // HYPRE_REAL_MIN
   double d2[2] = {0.0, 0.0};
   double alpha, beta, gamma, epsilon, temp, rres, r_norm, b_norm, den_norm;
   double r_norm_0, cf_ave_0 = 0.0, cf_ave_1 = 0.0;
   int iter = 0;
   const double *z = nullptr;
   memset(res, 0, sizeof(*res));

   // r0 = b - A x ; r = p = r0   (bicgstab.c:287-294)
   KRY_TRY(ops.matvec(-1.0, x, 1.0, b, r0));
   KRY_TRY(ops.copy(r0, r));
   KRY_TRY(ops.copy(r0, p));
   {
      const double *xs[2] = {b, r0};
      KRY_TRY(ops.dots(2, xs, xs, d2));
   }
   b_norm = sqrt(d2[0]);
   res->b_norm = b_norm;
   if (kry_not_finite(b_norm)) { res->error_flag = 1; return B200_OK; }
   rres = d2[1];
   r_norm = sqrt(rres);
   if (kry_not_finite(r_norm)) { res->error_flag = 1; return B200_OK; }
   r_norm_0 = r_norm;
   if (norms_h) { norms_h[0] = r_norm; }
   den_norm = b_norm > 0.0 ? b_norm : r_norm;
   epsilon = prm->a_tol > prm->tol * den_norm ? prm->a_tol : prm->tol * den_norm;
   if (prm->stop_crit) { epsilon = prm->a_tol == 0.0 ? prm->tol : prm->a_tol; }      // bicgstab.c:372-383
   if (b_norm > 0.0) { res->rel_residual_norm = r_norm / b_norm; }
   if (r_norm == 0.0) { return B200_OK; }                                             // :403-408
   if (r_norm <= epsilon && iter >= min_iter) { res->converged = 1; res->real_residual_exit = 1; return B200_OK; }  // :409-419

   while (iter < max_iter)
   {
      iter++;
      KRY_TRY(ops.precond(p, v, &z));
      KRY_TRY(ops.matvec(1.0, z, 0.0, nullptr, q));
      {
         const double *xs[1] = {r0}, *ys[1] = {q};
         KRY_TRY(ops.dots(1, xs, ys, d2));
      }
      temp = d2[0];
      if (fabs(temp) >= epsmac) { alpha = rres / temp; }
      else { res->error_flag |= 2; return B200_OK; }   // the reference returns with num_iterations still 0, :433-437
      KRY_TRY(ops.bicg_half(alpha, z, q, x, r));
      KRY_TRY(ops.precond(r, v, &z));
      KRY_TRY(ops.matvec(1.0, z, 0.0, nullptr, s));
      {
         const double *xs[2] = {r, s}, *ys[2] = {s, s};
         KRY_TRY(ops.dots(2, xs, ys, d2));
      }
      gamma = (d2[0] == 0.0 && d2[1] == 0.0) ? 0.0 : d2[0] / d2[1];
      KRY_TRY(ops.bicg_full(gamma, z, s, r0, x, r, d2));
      r_norm = sqrt(d2[0]);
      if (norms_h) { norms_h[iter] = r_norm; }
      if (r_norm <= epsilon && iter >= min_iter)                                      // :503-517
      {
         // the reference recomputes r = b - A x here and, when that residual is still above epsilon, goes on
         // with it; <r0,r> is then taken from the recomputed r
         KRY_TRY(ops.matvec(-1.0, x, 1.0, b, r));
         const double *xs[2] = {r, r0}, *ys[2] = {r, r};
         KRY_TRY(ops.dots(2, xs, ys, d2));
         r_norm = sqrt(d2[0]);
         if (r_norm <= epsilon) { res->converged = 1; res->real_residual_exit = 1; break; }
      }
      if (prm->cf_tol > 0.0 && kry_cf_exceeded(prm->cf_tol, r_norm, r_norm_0, iter, &cf_ave_0, &cf_ave_1)) { break; }
      if (fabs(rres) >= epsmac) { beta = 1.0 / rres; }
      else { res->error_flag |= 2; return B200_OK; }
      rres = d2[1];
      beta *= rres;
      if (!(fabs(gamma) >= epsmac)) { res->error_flag |= 2; return B200_OK; }
      KRY_TRY(ops.bicg_dir(gamma, beta * alpha / gamma, q, r, p));
   }
   res->num_iterations = iter;
   res->rel_residual_norm = b_norm > 0.0 ? r_norm / b_norm : r_norm;
   if (iter >= max_iter && r_norm > epsilon && epsilon > 0) { res->error_flag |= 256; }
   return B200_OK;
}

// LGMRES (hypre_LGMRESSolve, src/krylov/lgmres.c:337-...; ij -solver 51 / 50) with cf_tol = 0: restarted GMRES
// whose space is augmented with the last aug_dim error approximations.  p: k_dim + 1 basis vectors; av: aug_dim + 1
// vectors (the approximations, the last one a scratch slot); aav: aug_dim vectors (their images under A C);
// r, w: work vectors.  Same backend methods as gmres_core (modified Gram-Schmidt).
template <class B>
int lgmres_core(B &ops, const b200_lgmres_params *prm, const double *b, double *x, double *const *p,
                double *const *av, double *const *aav, double *r, double *w, b200_krylov_result *res, double *norms_h)
{
   const int k_dim = prm->k_dim, aug_dim = prm->aug_dim, kd = k_dim + aug_dim;
   const int max_iter = prm->max_iter, min_iter = prm->min_iter;
   const double epsmac = 1.e-16;
   std::vector<double> rs(kd + 1, 0.0), c(kd + 1, 0.0), s(kd + 1, 0.0), col(kd + 2, 0.0);
   std::vector<double> hh((size_t)(kd + 1) * (kd + 1), 0.0);
   std::vector<int> aug_order(aug_dim + 1, 0);
#define HH(a, bb) hh[(size_t)(a) * (kd + 1) + (bb)]
   double d2[2] = {0.0, 0.0};
   double b_norm, r_norm, r_norm_last = 0.0, den_norm, epsilon, t, gamma, tmp_norm;
   double r_norm_0, cf_ave_0 = 0.0, cf_ave_1 = 0.0;
   bool break_value = false;
   int i = 0, j, k, ii, iter = 0, aug_ct = 0, it_arnoldi, it_total, it_aug, order, spot = 0;
   const double *z = nullptr;
   memset(res, 0, sizeof(*res));

   KRY_TRY(ops.matvec(-1.0, x, 1.0, b, p[0]));
   {
      const double *xs[2] = {b, p[0]};
      KRY_TRY(ops.dots(2, xs, xs, d2));
   }
   b_norm = sqrt(d2[0]);
   res->b_norm = b_norm;
   if (kry_not_finite(b_norm)) { res->error_flag = 1; return B200_OK; }
   r_norm = sqrt(d2[1]);
   if (kry_not_finite(r_norm)) { res->error_flag = 1; return B200_OK; }
   r_norm_0 = r_norm;
   if (norms_h) { norms_h[0] = r_norm; }
   den_norm = b_norm > 0.0 ? b_norm : r_norm;
   epsilon = prm->a_tol > prm->tol * den_norm ? prm->a_tol : prm->tol * den_norm;

   while (iter < max_iter)
   {
      rs[0] = r_norm;
      if (r_norm == 0.0) { return B200_OK; }           // lgmres.c:546-560 (num_iterations stays 0)
      if (r_norm <= epsilon && iter >= min_iter)
      {
         KRY_TRY(ops.matvec(-1.0, x, 1.0, b, r));
         const double *xs[1] = {r};
         KRY_TRY(ops.dots(1, xs, xs, d2));
         r_norm = sqrt(d2[0]);
         if (r_norm <= epsilon) { res->real_residual_exit = 1; break; }
      }
      t = 1.0 / r_norm;
      r_norm_last = r_norm;
      KRY_TRY(ops.scale(t, p[0]));
      i = 0;
      // number of Arnoldi steps of this cycle: the space keeps k_dim vectors (approx_constant) (:588-603)
      it_arnoldi = prm->approx_constant ? k_dim - aug_ct : k_dim - aug_dim;
      it_total = it_arnoldi + aug_ct;
      it_aug = 0;
      while (i < it_total && iter < max_iter)
      {
         i++; iter++;
         if (i <= it_arnoldi)
         {
            KRY_TRY(ops.precond(p[i - 1], r, &z));
            KRY_TRY(ops.matvec(1.0, z, 0.0, nullptr, p[i]));
         }
         else
         {
            // augmentation step: the stored image of the approximation of this order (:617-636)
            it_aug++;
            order = i - it_arnoldi - 1;
            for (ii = 0; ii < aug_dim; ii++) { if (aug_order[ii] == order) { spot = ii; break; } }
            KRY_TRY(ops.copy(aav[spot], p[i]));
         }
         KRY_TRY(ops.mgs(i, p, col.data()));
         for (j = 0; j < i; j++) { HH(j, i - 1) = col[j]; }
         t = sqrt(col[i]);
         HH(i, i - 1) = t;
         if (t != 0.0) { t = 1.0 / t; KRY_TRY(ops.scale(t, p[i])); }
         for (j = 1; j < i; j++)
         {
            t = HH(j - 1, i - 1);
            HH(j - 1, i - 1) = s[j - 1] * HH(j, i - 1) + c[j - 1] * t;
            HH(j, i - 1) = -s[j - 1] * t + c[j - 1] * HH(j, i - 1);
         }
         t = HH(i, i - 1) * HH(i, i - 1);
         t += HH(i - 1, i - 1) * HH(i - 1, i - 1);
         gamma = sqrt(t);
         if (gamma == 0.0) { gamma = epsmac; }
         c[i - 1] = HH(i - 1, i - 1) / gamma;
         s[i - 1] = HH(i, i - 1) / gamma;
         rs[i] = -HH(i, i - 1) * rs[i - 1];
         rs[i] /= gamma;
         rs[i - 1] = c[i - 1] * rs[i - 1];
         HH(i - 1, i - 1) = s[i - 1] * HH(i, i - 1) + c[i - 1] * HH(i - 1, i - 1);
         r_norm = fabs(rs[i]);
         if (norms_h) { norms_h[iter] = r_norm; }
         if (prm->cf_tol > 0.0 && kry_cf_exceeded(prm->cf_tol, r_norm, r_norm_0, iter, &cf_ave_0, &cf_ave_1)) { break_value = true; break; }
         if (r_norm <= epsilon && iter >= min_iter) { break; }
      }
      if (break_value) { break; }
      rs[i - 1] = rs[i - 1] / HH(i - 1, i - 1);
      for (k = i - 2; k >= 0; k--)
      {
         t = 0.0;
         for (j = k + 1; j < i; j++) { t -= HH(k, j) * rs[j]; }
         t += rs[k];
         rs[k] = t / HH(k, k);
      }
      // correction before the preconditioner = combination of the Arnoldi directions and of the approximations
      // that served in this cycle (:712-755)
      if (it_arnoldi > i) { it_arnoldi = i; }
      if (!it_aug)
      {
         KRY_TRY(ops.copy(p[i - 1], w));
         KRY_TRY(ops.scale(rs[i - 1], w));
         for (j = i - 2; j >= 0; j--) { KRY_TRY(ops.axpy(rs[j], p[j], w)); }
      }
      else
      {
         KRY_TRY(ops.copy(p[0], w));
         KRY_TRY(ops.scale(rs[0], w));
         for (j = 1; j < it_arnoldi; j++) { KRY_TRY(ops.axpy(rs[j], p[j], w)); }
         for (ii = 0; ii < it_aug; ii++)
         {
            for (j = 0; j < aug_dim; j++) { if (aug_order[j] == ii) { spot = j; break; } }
            KRY_TRY(ops.axpy(rs[it_arnoldi + ii], av[spot], w));
         }
      }
      KRY_TRY(ops.copy(w, av[aug_dim]));          // the new approximation, kept aside until the cycle is closed
      KRY_TRY(ops.precond(w, r, &z));
      KRY_TRY(ops.axpy(1.0, z, x));

      if (r_norm <= epsilon && iter >= min_iter)
      {
         KRY_TRY(ops.matvec(-1.0, x, 1.0, b, r));
         const double *xs[1] = {r};
         KRY_TRY(ops.dots(1, xs, xs, d2));
         r_norm = sqrt(d2[0]);
         if (r_norm <= epsilon) { res->converged = 1; res->real_residual_exit = 1; break; }
         KRY_TRY(ops.copy(r, p[0]));
         i = 0;
      }

      // r0 (unscaled) into w; residual vector of the restart into p[0] (:785-810)
      KRY_TRY(ops.copy(p[0], w));
      KRY_TRY(ops.scale(r_norm_last, w));
      for (j = i; j > 0; j--)
      {
         rs[j - 1] = -s[j - 1] * rs[j];
         rs[j] = c[j - 1] * rs[j];
      }
      if (i) { KRY_TRY(ops.axpy(rs[i] - 1.0, p[i], p[i])); }
      for (j = i - 1; j > 0; j--) { KRY_TRY(ops.axpy(rs[j], p[j], p[i])); }
      if (i)
      {
         KRY_TRY(ops.axpy(rs[0] - 1.0, p[0], p[0]));
         KRY_TRY(ops.axpy(1.0, p[i], p[0]));
      }

      // keep the new approximation (normalised) and its image A C z = (r0 - rm) / ||z|| for later cycles (:812-860)
      if (aug_dim > 0)
      {
         if (!aug_ct) { spot = 0; aug_ct++; }
         else if (aug_ct < aug_dim) { spot = aug_ct; aug_ct++; }
         else { for (ii = 0; ii < aug_dim; ii++) { if (aug_order[ii] == aug_dim - 1) { spot = ii; } } }
         KRY_TRY(ops.copy(av[aug_dim], av[spot]));
         const double *xs[1] = {av[spot]};
         KRY_TRY(ops.dots(1, xs, xs, d2));
         tmp_norm = 1.0 / sqrt(d2[0]);
         KRY_TRY(ops.scale(tmp_norm, av[spot]));
         for (ii = 0; ii < aug_dim; ii++) { aug_order[ii]++; }
         aug_order[spot] = 0;
         KRY_TRY(ops.copy(w, aav[spot]));
         KRY_TRY(ops.scale(-1.0, aav[spot]));
         KRY_TRY(ops.axpy(1.0, p[0], aav[spot]));
         KRY_TRY(ops.scale(-tmp_norm, aav[spot]));
      }
   }
#undef HH
   res->num_iterations = iter;
   res->rel_residual_norm = b_norm > 0.0 ? r_norm / b_norm : r_norm;
   if (iter >= max_iter && r_norm > epsilon && epsilon > 0) { res->error_flag |= 256; }
   return B200_OK;
}

