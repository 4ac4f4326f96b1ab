/* amg_oracle_cycle.c -- TEST INFRASTRUCTURE: plain-C restatement of
 * hypre_BoomerAMGCycle (parcsr_ls/par_cycle.c:22-803) and hypre_PCGSolve
 * (krylov/pcg.c:317-1008) for one rank, on borrowed CSR arrays. */
#include "amg_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
   int n, n_coarse;
   const int *Ai, *Aj, *Pi, *Pj;
   const double *Aa, *Pa, *l1, *ds, *coefs;
   double relax_weight;
   double *F, *U;          /* owned for level > 0 */
   int all_zeros;          /* hypre_ParVectorAllZeros(U) */
   int nblocks; int *starts;   /* N-rank row partition (two-stage GS only); NULL = one rank */
   const int *cf;          /* CF_marker_array[level] (borrowed), NULL on the coarsest level */
} orc_level;

struct orc_amg {
   int num_levels;
   orc_level *L;
   int relax_type[4], num_sweeps[4], cycle_type, fcycle, cheby_order, cheby_scale;
   int coarse_n;
   int precond_cycles;     /* hypre_ParAMGDataMaxIter of a preconditioner (ij -precon_cycles), default 1 */
   int relax_order;        /* hypre_ParAMGDataRelaxOrder: 1 = C points then F points (down), F then C (up) */
   double *A_mat, *A_work, *b_vec;
   double *Vtemp, *Ztemp, *Ptemp, *Rtemp;
   int maxn;
};

orc_amg *orc_amg_create(int num_levels)
{
   orc_amg *h = (orc_amg *) calloc(1, sizeof(orc_amg));
   h->num_levels = num_levels;
   h->L = (orc_level *) calloc((size_t) num_levels, sizeof(orc_level));
   h->relax_type[1] = h->relax_type[2] = 18; h->relax_type[3] = 9;
   h->num_sweeps[0] = h->num_sweeps[1] = h->num_sweeps[2] = h->num_sweeps[3] = 1;
   h->cycle_type = 1; h->cheby_order = 2; h->cheby_scale = 1;
   h->precond_cycles = 1;
   return h;
}

void orc_amg_destroy(orc_amg *h)
{
   int l;
   if (!h) { return; }
   for (l = 1; l < h->num_levels; l++) { free(h->L[l].F); free(h->L[l].U); }
   for (l = 0; l < h->num_levels; l++) { free(h->L[l].starts); }
   free(h->L); free(h->A_mat); free(h->A_work); free(h->b_vec);
   free(h->Vtemp); free(h->Ztemp); free(h->Ptemp); free(h->Rtemp);
   free(h);
}

void orc_amg_set_level(orc_amg *h, int level, int n, const int *Ai, const int *Aj, const double *Aa,
                       int n_coarse, const int *Pi, const int *Pj, const double *Pa,
                       const double *l1, const double *ds, const double *coefs, double relax_weight)
{
   orc_level *L = &h->L[level];
   L->n = n; L->Ai = Ai; L->Aj = Aj; L->Aa = Aa;
   L->n_coarse = n_coarse; L->Pi = Pi; L->Pj = Pj; L->Pa = Pa;
   L->l1 = l1; L->ds = ds; L->coefs = coefs; L->relax_weight = relax_weight;
   if (level > 0)
   {
      free(L->F); free(L->U);
      L->F = (double *) calloc((size_t) (n > 0 ? n : 1), sizeof(double));
      L->U = (double *) calloc((size_t) (n > 0 ? n : 1), sizeof(double));
   }
   if (n > h->maxn)
   {
      h->maxn = n;
      free(h->Vtemp); free(h->Ztemp); free(h->Ptemp); free(h->Rtemp);
      h->Vtemp = (double *) calloc((size_t) n, sizeof(double));
      h->Ztemp = (double *) calloc((size_t) n, sizeof(double));
      h->Ptemp = (double *) calloc((size_t) n, sizeof(double));
      h->Rtemp = (double *) calloc((size_t) n, sizeof(double));
   }
}

void orc_amg_set_cycle(orc_amg *h, const int relax_type[4], const int num_sweeps[4], int cycle_type,
                       int fcycle, int cheby_order, int cheby_scale)
{
   int k;
   for (k = 0; k < 4; k++) { h->relax_type[k] = relax_type[k]; h->num_sweeps[k] = num_sweeps[k]; }
   h->cycle_type = cycle_type; h->fcycle = fcycle;
   h->cheby_order = cheby_order; h->cheby_scale = cheby_scale;
}

void orc_amg_set_level_blocks(orc_amg *h, int level, int nblocks, const int *starts)
{
   orc_level *L = &h->L[level];
   int k;
   free(L->starts);
   L->starts = NULL; L->nblocks = 0;
   if (nblocks > 1 && starts)
   {
      L->starts = (int *) malloc(sizeof(int) * (size_t)(nblocks + 1));
      for (k = 0; k <= nblocks; k++) { L->starts[k] = starts[k]; }
      L->nblocks = nblocks;
   }
}

void orc_amg_set_coarse_dense(orc_amg *h, int n, const double *A_mat)
{
   free(h->A_mat); free(h->A_work); free(h->b_vec);
   h->coarse_n = n;
   h->A_mat  = (double *) malloc(sizeof(double) * (size_t) n * n);
   h->A_work = (double *) malloc(sizeof(double) * (size_t) n * n);
   h->b_vec  = (double *) malloc(sizeof(double) * (size_t) n);
   memcpy(h->A_mat, A_mat, sizeof(double) * (size_t) n * n);
}

void orc_amg_set_level_cf(orc_amg *h, int level, const int *cf_marker) { h->L[level].cf = cf_marker; }
void orc_amg_set_relax_order(orc_amg *h, int relax_order) { h->relax_order = relax_order; }

/* one call of hypre_BoomerAMGRelax on `level` (par_relax.c:23-175; par_cycle.c:395-612 dispatch).
 * points: 0 all rows, 1 / -1 only the rows whose CF marker has that value (relax types 0, 7, 18; the two-stage
 * Gauss-Seidel sweeps ignore it, par_relax.c:1597-1607) */
static int orc_relax(orc_amg *h, int level, int relax_type, int points, const double *f, double *u)
{
   orc_level *L = &h->L[level];
   int err = 0;
   switch (relax_type)
   {
      case 18: case 7:
         if (points && L->cf)
         {
            /* hypre_BoomerAMGRelax7Jacobi (par_relax.c:1177-1254): the whole Vtemp = w (f - A u), then
             * hypre_ParVectorElmdivpyMarked (seq_mv/vector.c:823-870): u_i += Vtemp_i / l1_i where marker_i == points */
            int i;
            if (L->all_zeros) { for (i = 0; i < L->n; i++) { h->Vtemp[i] = L->relax_weight * f[i]; } }
            else { orc_csr_matvec(L->n, L->Ai, L->Aj, L->Aa, -L->relax_weight, u, L->relax_weight, f, h->Vtemp); }
            for (i = 0; i < L->n; i++) { if (L->cf[i] == points) { u[i] += h->Vtemp[i] / L->l1[i]; } }
            break;
         }
         orc_relax_l1jacobi(L->n, L->Ai, L->Aj, L->Aa, f, L->l1, L->relax_weight, u, h->Vtemp, L->all_zeros);
         break;
      case 0:
      {
         /* hypre_BoomerAMGRelax0WeightedJacobi (par_relax.c:320-335) -> hypre_BoomerAMGRelaxWeightedJacobi_core
          * with Skip_diag = 1 and no l1 norms (:179-314): u_i = (1-w) u_i + w (f_i - sum_{j != i} a_ij u_j) / a_ii,
          * the diagonal entry first in its row; rows with a zero diagonal are skipped */
         const double w = L->relax_weight, omw = 1.0 - L->relax_weight;
         int i, jj;
         memcpy(h->Vtemp, u, sizeof(double) * (size_t) L->n);
         for (i = 0; i < L->n; i++)
         {
            const double di = L->Aa[L->Ai[i]];
            if ((points == 0 || L->cf == NULL || L->cf[i] == points) && di != 0.0)
            {
               double res = f[i];
               for (jj = L->Ai[i] + 1; jj < L->Ai[i + 1]; jj++) { res -= L->Aa[jj] * h->Vtemp[L->Aj[jj]]; }
               u[i] *= omw;
               u[i] += w * res / di;
            }
         }
         break;
      }
      case 16:
         /* hypre_ParCSRRelax_Cheby_Solve(A, f, ds, coefs, order, scale, variant, u, Vtemp, Ztemp, Ptemp, Rtemp) */
         orc_relax_cheby(L->n, L->Ai, L->Aj, L->Aa, f, L->ds, L->coefs, h->cheby_order, h->cheby_scale,
                         u, h->Vtemp, h->Ztemp, h->Ptemp, h->Rtemp);
         break;
      case 11: case 12:
         err = orc_relax_gs2_blocks(L->n, L->Ai, L->Aj, L->Aa, f, L->relax_weight, relax_type == 11 ? 1 : 2, u, h->Vtemp,
                                    L->nblocks, L->starts);
         break;
      case 9:
         /* hypre_GaussElimSolve type 9 (par_gauss_elim.c:563-651): copy A_mat, eliminate, copy slice */
         memcpy(h->A_work, h->A_mat, sizeof(double) * (size_t) h->coarse_n * h->coarse_n);
         memcpy(h->b_vec, f, sizeof(double) * (size_t) h->coarse_n);
         err = orc_gselim(h->A_work, h->b_vec, h->coarse_n);
         memcpy(u, h->b_vec, sizeof(double) * (size_t) h->coarse_n);
         break;
      default:
         return -1;
   }
   L->all_zeros = 0;   /* par_relax.c:170 */
   return err;
}

/* hypre_BoomerAMGCycle: level-counter state machine of par_cycle.c:209-223,647-779 */
int orc_amg_cycle(orc_amg *h, const double *f, double *u, int all_zeros)
{
   const int nl = h->num_levels;
   int *lev_counter = (int *) calloc((size_t) nl, sizeof(int));
   int k, j, level = 0, cycle_param = 1, not_finished = 1, fcycle_lev = nl - 2, err = 0;
   lev_counter[0] = 1;
   for (k = 1; k < nl; k++) { lev_counter[k] = h->fcycle ? 1 : h->cycle_type; }
   h->L[0].all_zeros = all_zeros;
   while (not_finished)
   {
      const double *F = level == 0 ? f : h->L[level].F;
      double *U = level == 0 ? u : h->L[level].U;
      int relax_type = nl > 1 ? h->relax_type[cycle_param] : h->relax_type[0];
      int num_sweep = nl > 1 ? h->num_sweeps[cycle_param] : h->num_sweeps[0];
      for (j = 0; j < num_sweep; j++)
      {
         /* hypre_BoomerAMGRelaxIF (par_relax_interface.c:20-64): with relax_order 1 a sweep on a non-coarsest level is
          * two calls, C then F points on the way down, F then C on the way up; Chebyshev (16) and Gaussian
          * elimination (9) do not go through it (par_cycle.c:495-540) */
         const int cf_order = h->relax_order == 1 && cycle_param < 3 && nl > 1 && relax_type != 16 && relax_type != 9;
         int pass;
         for (pass = 0; pass < (cf_order ? 2 : 1); pass++)
         {
            const int points = !cf_order ? 0 : ((cycle_param < 2) == (pass == 0) ? 1 : -1);
            int e = orc_relax(h, level, relax_type, points, F, U);
            if (e < 0) { free(lev_counter); return e; }
            err |= e;
         }
      }
      --lev_counter[level];
      if (lev_counter[level] >= 0 && level != nl - 1)
      {
         orc_level *L = &h->L[level], *C = &h->L[level + 1];
         memset(C->U, 0, sizeof(double) * (size_t) C->n);   /* hypre_ParVectorSetZeros */
         C->all_zeros = 1;
         orc_csr_matvec(L->n, L->Ai, L->Aj, L->Aa, -1.0, U, 1.0, F, h->Vtemp);          /* residual   */
         orc_csr_matvecT(L->n, C->n, L->Pi, L->Pj, L->Pa, 1.0, h->Vtemp, 0.0, C->F);    /* restriction */
         ++level;
         lev_counter[level] = lev_counter[level] > h->cycle_type ? lev_counter[level] : h->cycle_type;
         cycle_param = (level == nl - 1) ? 3 : 1;
      }
      else if (level != 0)
      {
         orc_level *Lf = &h->L[level - 1];
         double *Uf = level - 1 == 0 ? u : Lf->U;
         orc_csr_matvec(Lf->n, Lf->Pi, Lf->Pj, Lf->Pa, 1.0, h->L[level].U, 1.0, Uf, Uf);  /* interpolation */
         Lf->all_zeros = 0;
         --level;
         cycle_param = 2;
         if (h->fcycle && fcycle_lev == level)
         {
            lev_counter[level] = lev_counter[level] > 1 ? lev_counter[level] : 1;
            fcycle_lev--;
         }
      }
      else { not_finished = 0; }
   }
   free(lev_counter);
   return err;
}

void orc_amg_set_precond_cycles(orc_amg *h, int cycles) { h->precond_cycles = cycles < 1 ? 1 : cycles; }

/* BoomerAMG as a preconditioner: hypre_BoomerAMGSolve with tol 0 and max_iter k runs k cycles without looking at
 * the residual (par_amg_solve.c:257-336; tol > 0 would compute norms); the Krylov solver cleared z before the call
 * (ClearVector sets all_zeros), so the first cycle starts from a zero guess and the others continue on it. */
static void orc_amg_precond(orc_amg *h, const double *r, double *z)
{
   int k;
   memset(z, 0, sizeof(double) * (size_t) h->L[0].n);
   orc_amg_cycle(h, r, z, 1);
   for (k = 1; k < h->precond_cycles; k++) { orc_amg_cycle(h, r, z, 0); }
}

static double orc_dot(int n, const double *x, const double *y)
{
   double s = 0.0;
   int i;
   for (i = 0; i < n; i++) { s += x[i] * y[i]; }
   return s;
}

/* hypre_PCGSolve with the default stopping rule (pcg.c:447-470,563-977); h == level-0 operator +
 * AMG preconditioner (one cycle, zero initial guess: ClearVector sets all_zeros, par_krylov_func.c:232-236) */
int orc_pcg(orc_amg *h, const double *b, double *x, double tol, int max_iter, int two_norm,
            double *norms, double *rel_norms, int *converged)
{
   orc_level *L0 = &h->L[0];
   const int n = L0->n;
   double *p = (double *) calloc((size_t) n, sizeof(double));
   double *s = (double *) calloc((size_t) n, sizeof(double));
   double *r = (double *) calloc((size_t) n, sizeof(double));
   double bi_prod, eps, gamma, gamma_old, alpha, beta, sdotp, i_prod = 0.0;
   int i = 0, k;
   if (converged) { *converged = 0; }
   if (two_norm) { bi_prod = orc_dot(n, b, b); }
   else { orc_amg_precond(h, b, p); bi_prod = orc_dot(n, p, b); }
   eps = tol * tol;
   if (bi_prod <= 0.0) { memcpy(x, b, sizeof(double) * (size_t) n); free(p); free(s); free(r); return 0; }
   orc_csr_matvec(n, L0->Ai, L0->Aj, L0->Aa, -1.0, x, 1.0, b, r);
   orc_amg_precond(h, r, p);
   gamma = orc_dot(n, r, p);
   if (norms) { norms[0] = sqrt(two_norm ? orc_dot(n, r, r) : gamma); }
   while ((i + 1) <= max_iter)
   {
      i++;
      orc_csr_matvec(n, L0->Ai, L0->Aj, L0->Aa, 1.0, p, 0.0, NULL, s);
      sdotp = orc_dot(n, s, p);
      if (sdotp == 0.0) { break; }
      alpha = gamma / sdotp;
      if (alpha <= 0.0) { break; }
      gamma_old = gamma;
      for (k = 0; k < n; k++) { x[k] += alpha * p[k]; }
      for (k = 0; k < n; k++) { r[k] += -alpha * s[k]; }
      orc_amg_precond(h, r, s);
      gamma = orc_dot(n, r, s);
      i_prod = two_norm ? orc_dot(n, r, r) : gamma;
      if (norms) { norms[i] = sqrt(i_prod); }
      if (rel_norms) { rel_norms[i] = sqrt(i_prod / bi_prod); }
      if (i_prod / bi_prod < eps) { if (converged) { *converged = 1; } break; }
      if (gamma <= 0.0) { break; }
      beta = gamma / gamma_old;
      for (k = 0; k < n; k++) { p[k] = beta * p[k]; }
      for (k = 0; k < n; k++) { p[k] += s[k]; }
   }
   free(p); free(s); free(r);
   return i;
}

/* Options the Krylov restatements below read at the start of a solve (test infrastructure: set, solve, reset):
 * cf_tol = convergence-factor tolerance (gmres.c:632-652, bicgstab.c:519-540, ...), stop_crit = BiCGSTAB's absolute
 * stopping rule (bicgstab.c:372-392). */
static double g_orc_cf_tol = 0.0;
static int    g_orc_stop_crit = 0;
void orc_krylov_set_options(double cf_tol, int stop_crit) { g_orc_cf_tol = cf_tol; g_orc_stop_crit = stop_crit; }

/* the reference's "adequate progress" test: returns 1 when the weighted average convergence factor exceeds cf_tol */
static int orc_cf_exceeded(double cf_tol, double r_norm, double r_norm_0, int iter, double *cf_ave_0, double *cf_ave_1)
{
   double weight;
   *cf_ave_0 = *cf_ave_1;
   *cf_ave_1 = pow(r_norm / r_norm_0, 1.0 / (2.0 * iter));
   weight = fabs(*cf_ave_1 - *cf_ave_0);
   weight = weight / (*cf_ave_1 > *cf_ave_0 ? *cf_ave_1 : *cf_ave_0);
   weight = 1.0 - weight;
   return weight * *cf_ave_1 > cf_tol;
}

/* ---- preconditioner shared by the GMRES / BiCGSTAB restatements below.
 * kind 0: one AMG cycle with a zero initial guess (the callers ClearVector first, which sets all_zeros);
 * kind 1: HYPRE_ParCSRDiagScale (parcsr_ls/HYPRE_parcsr_pcg.c:267-277): z = r ./ a_ii, diagonal entry first
 * in its row; kind 2: hypre_ParKrylovIdentity (par_krylov_func.c:312-330): z = r. */
static void orc_precond(orc_amg *h, int kind, const double *r, double *z)
{
   orc_level *L0 = &h->L[0];
   const int n = L0->n;
   int k;
   if (kind == 0) { orc_amg_precond(h, r, z); }
   else if (kind == 1) { for (k = 0; k < n; k++) { z[k] = r[k] / L0->Aa[L0->Ai[k]]; } }
   else { memcpy(z, r, sizeof(double) * (size_t) n); }
}

/* hypre_GMRESSolve (krylov/gmres.c:274-930) with rel_change = 0 and cf_tol = 0 (what ij -solver 3/4 passes):
 * right-preconditioned restarted GMRES(k_dim), modified Gram-Schmidt, Givens rotations, true-residual
 * check on convergence.  norms[0..iters] = the residual norms the reference logs.  Returns the iteration
 * count; *converged as gmres_data->converged; *rel_resid as gmres_data->rel_residual_norm. */
static int orc_gmres_impl(orc_amg *h, int precond_kind, const double *b, double *x, double tol, double a_tol, int k_dim,
                          int min_iter, int max_iter, int skip_real_r_check, double *norms, double *rel_resid,
                          int *converged, int cgs, int flex)
{
   orc_level *L0 = &h->L[0];
   const int n = L0->n;
   const double epsmac = 1.e-16;
   double **p = (double **) calloc((size_t) k_dim + 1, sizeof(double *));
   double *r = (double *) calloc((size_t) n, sizeof(double));
   double *w = (double *) calloc((size_t) n, sizeof(double));
   double *rs = (double *) calloc((size_t) k_dim + 1, sizeof(double));
   double *c = (double *) calloc((size_t) k_dim, sizeof(double));
   double *s = (double *) calloc((size_t) k_dim, sizeof(double));
   double **hh = (double **) calloc((size_t) k_dim + 1, sizeof(double *));
   double **pv = (double **) calloc((size_t) k_dim + 1, sizeof(double *));   /* flexgmres.c: pre_vecs */
   double b_norm, r_norm, den_norm, epsilon, t, gamma, real_r_norm_old, real_r_norm_new;
   const double cf_tol = g_orc_cf_tol;
   double r_norm_0, cf_ave_0 = 0.0, cf_ave_1 = 0.0;
   int i, j, k, q, iter = 0, break_value = 0;
   for (i = 0; i <= k_dim; i++)
   {
      p[i] = (double *) calloc((size_t) n, sizeof(double));
      hh[i] = (double *) calloc((size_t) k_dim, sizeof(double));
      pv[i] = flex ? (double *) calloc((size_t) n, sizeof(double)) : NULL;
   }
   if (converged) { *converged = 0; }
   orc_csr_matvec(n, L0->Ai, L0->Aj, L0->Aa, -1.0, x, 1.0, b, p[0]);
   b_norm = sqrt(orc_dot(n, b, b));
   real_r_norm_old = b_norm;
   r_norm = sqrt(orc_dot(n, p[0], p[0]));
   r_norm_0 = r_norm;
   if (norms) { norms[0] = r_norm; }
   den_norm = b_norm > 0.0 ? b_norm : r_norm;
   epsilon = a_tol > tol * den_norm ? a_tol : tol * den_norm;
   while (iter < max_iter)
   {
      rs[0] = r_norm;
      if (r_norm == 0.0) { goto finish_no_norm; }
      if (r_norm <= epsilon && iter >= min_iter)
      {
         orc_csr_matvec(n, L0->Ai, L0->Aj, L0->Aa, -1.0, x, 1.0, b, r);
         r_norm = sqrt(orc_dot(n, r, r));
         if (r_norm <= epsilon) { break; }
      }
      t = 1.0 / r_norm;
      for (q = 0; q < n; q++) { p[0][q] *= t; }
      i = 0;
      while (i < k_dim && iter < max_iter)
      {
         i++; iter++;
         if (flex)
         {
            orc_precond(h, precond_kind, p[i - 1], pv[i - 1]);
            orc_csr_matvec(n, L0->Ai, L0->Aj, L0->Aa, 1.0, pv[i - 1], 0.0, NULL, p[i]);
         }
         else
         {
            orc_precond(h, precond_kind, p[i - 1], r);
            orc_csr_matvec(n, L0->Ai, L0->Aj, L0->Aa, 1.0, r, 0.0, NULL, p[i]);
         }
         if (cgs)
         {
            /* hypre_COGMRESSolve with cgs == 1 (cogmres.c:553-577): all inner products from the same p[i]
             * (MassInnerProd), then one multi-vector update (MassAxpy, vector_batched.c:212-247) */
            for (j = 0; j < i; j++) { hh[j][i - 1] = orc_dot(n, p[i], p[j]); }
            for (j = 0; j < i; j++) { for (q = 0; q < n; q++) { p[i][q] += -hh[j][i - 1] * p[j][q]; } }
         }
         else
         {
            for (j = 0; j < i; j++)
            {
               hh[j][i - 1] = orc_dot(n, p[j], p[i]);
               for (q = 0; q < n; q++) { p[i][q] += -hh[j][i - 1] * p[j][q]; }
            }
         }
         t = sqrt(orc_dot(n, p[i], p[i]));
         hh[i][i - 1] = t;
         if (t != 0.0) { t = 1.0 / t; for (q = 0; q < n; q++) { p[i][q] *= t; } }
         for (j = 1; j < i; j++)
         {
            t = hh[j - 1][i - 1];
            hh[j - 1][i - 1] = s[j - 1] * hh[j][i - 1] + c[j - 1] * t;
            hh[j][i - 1] = -s[j - 1] * t + c[j - 1] * hh[j][i - 1];
         }
         t = hh[i][i - 1] * hh[i][i - 1];
         t += hh[i - 1][i - 1] * hh[i - 1][i - 1];
         gamma = sqrt(t);
         if (gamma == 0.0) { gamma = epsmac; }
         c[i - 1] = hh[i - 1][i - 1] / gamma;
         s[i - 1] = hh[i][i - 1] / gamma;
         rs[i] = -hh[i][i - 1] * rs[i - 1];
         rs[i] /= gamma;
         rs[i - 1] = c[i - 1] * rs[i - 1];
         hh[i - 1][i - 1] = s[i - 1] * hh[i][i - 1] + c[i - 1] * hh[i - 1][i - 1];
         r_norm = fabs(rs[i]);
         if (norms) { norms[iter] = r_norm; }
         if (cf_tol > 0.0 && orc_cf_exceeded(cf_tol, r_norm, r_norm_0, iter, &cf_ave_0, &cf_ave_1)) { break_value = 1; break; }
         if (r_norm <= epsilon && iter >= min_iter) { break; }
      }
      if (break_value) { break; }        /* cf_tol: leave without the update of this cycle (gmres.c:732-735) */
      /* solution of the upper triangular system, update of x through the preconditioner */
      rs[i - 1] = rs[i - 1] / hh[i - 1][i - 1];
      for (k = i - 2; k >= 0; k--)
      {
         t = 0.0;
         for (j = k + 1; j < i; j++) { t -= hh[k][j] * rs[j]; }
         t += rs[k];
         rs[k] = t / hh[k][k];
      }
      if (flex)
      {
         /* hypre_FlexGMRESSolve (flexgmres.c:575-592): the correction is a combination of the stored
          * preconditioned vectors */
         memcpy(w, pv[i - 1], sizeof(double) * (size_t) n);
         for (q = 0; q < n; q++) { w[q] *= rs[i - 1]; }
         for (j = i - 2; j >= 0; j--) { for (q = 0; q < n; q++) { w[q] += rs[j] * pv[j][q]; } }
         for (q = 0; q < n; q++) { x[q] += w[q]; }
      }
      else
      {
         memcpy(w, p[i - 1], sizeof(double) * (size_t) n);
         for (q = 0; q < n; q++) { w[q] *= rs[i - 1]; }
         for (j = i - 2; j >= 0; j--) { for (q = 0; q < n; q++) { w[q] += rs[j] * p[j][q]; } }
         orc_precond(h, precond_kind, w, r);
         for (q = 0; q < n; q++) { x[q] += r[q]; }
      }
      if (r_norm <= epsilon && iter >= min_iter)
      {
         if (skip_real_r_check && !flex) { if (converged) { *converged = 1; } break; }
         orc_csr_matvec(n, L0->Ai, L0->Aj, L0->Aa, -1.0, x, 1.0, b, r);
         real_r_norm_new = r_norm = sqrt(orc_dot(n, r, r));
         if (r_norm <= epsilon) { if (converged) { *converged = 1; } break; }
         /* flexgmres.c:598-618 has no "residual no longer decreases" exit */
         if (!flex && real_r_norm_new >= real_r_norm_old) { if (converged) { *converged = 1; } break; }
         memcpy(p[0], r, sizeof(double) * (size_t) n);
         i = 0;
         real_r_norm_old = real_r_norm_new;
      }
      /* residual vector of the restart */
      for (j = i; j > 0; j--)
      {
         rs[j - 1] = -s[j - 1] * rs[j];
         rs[j] = c[j - 1] * rs[j];
      }
      if (i) { for (q = 0; q < n; q++) { p[i][q] += (rs[i] - 1.0) * p[i][q]; } }
      for (j = i - 1; j > 0; j--) { for (q = 0; q < n; q++) { p[i][q] += rs[j] * p[j][q]; } }
      if (i)
      {
         for (q = 0; q < n; q++) { p[0][q] += (rs[0] - 1.0) * p[0][q]; }
         for (q = 0; q < n; q++) { p[0][q] += p[i][q]; }
      }
   }
   if (rel_resid) { *rel_resid = b_norm > 0.0 ? r_norm / b_norm : r_norm; }
finish_no_norm:
   for (i = 0; i <= k_dim; i++) { free(p[i]); free(hh[i]); free(pv[i]); }
   free(p); free(hh); free(pv); free(r); free(w); free(rs); free(c); free(s);
   return iter;
}

int orc_gmres(orc_amg *h, int precond_kind, const double *b, double *x, double tol, double a_tol, int k_dim,
              int min_iter, int max_iter, int skip_real_r_check, double *norms, double *rel_resid, int *converged)
{
   return orc_gmres_impl(h, precond_kind, b, x, tol, a_tol, k_dim, min_iter, max_iter, skip_real_r_check, norms,
                         rel_resid, converged, 0, 0);
}

/* hypre_FlexGMRESSolve (krylov/flexgmres.c) with the default modify_pc (ij -solver 61 AMG / 60 diagonal scaling) */
int orc_flexgmres(orc_amg *h, int precond_kind, const double *b, double *x, double tol, double a_tol, int k_dim,
                  int min_iter, int max_iter, int skip_real_r_check, double *norms, double *rel_resid, int *converged)
{
   return orc_gmres_impl(h, precond_kind, b, x, tol, a_tol, k_dim, min_iter, max_iter, skip_real_r_check, norms,
                         rel_resid, converged, 0, 1);
}

/* hypre_COGMRESSolve (krylov/cogmres.c) with cgs = 1, rel_change = 0, cf_tol = 0 (ij -solver 16/17): GMRES with
 * classical Gram-Schmidt through MassInnerProd / MassAxpy; everything else as hypre_GMRESSolve */
int orc_cogmres(orc_amg *h, int precond_kind, const double *b, double *x, double tol, double a_tol, int k_dim,
                int min_iter, int max_iter, int skip_real_r_check, double *norms, double *rel_resid, int *converged)
{
   return orc_gmres_impl(h, precond_kind, b, x, tol, a_tol, k_dim, min_iter, max_iter, skip_real_r_check, norms,
                         rel_resid, converged, 1, 0);
}

/* hypre_BiCGSTABSolve (krylov/bicgstab.c:226-640) with stop_crit = 0 and cf_tol = 0 (ij -solver 9/10).
 * Returns the iteration count, or -(iter) - 1 when the reference would report a breakdown
 * (hypre_error_w_msg "BiCGSTAB broke down", :433-437, :519-523, :533-538). */
int orc_bicgstab(orc_amg *h, int precond_kind, const double *b, double *x, double tol, double a_tol,
                 int min_iter, int max_iter, double *norms, double *rel_resid, int *converged)
{
   orc_level *L0 = &h->L[0];
   const int n = L0->n;
   const double epsmac = 2.2250738585072014e-308;   /* HYPRE_REAL_MIN */
   double *r = (double *) calloc((size_t) n, sizeof(double));
   double *r0 = (double *) calloc((size_t) n, sizeof(double));
   double *s = (double *) calloc((size_t) n, sizeof(double));
   double *v = (double *) calloc((size_t) n, sizeof(double));
   double *p = (double *) calloc((size_t) n, sizeof(double));
   double *q = (double *) calloc((size_t) n, sizeof(double));
   double alpha, beta, gamma, epsilon, temp, res, r_norm, b_norm, den_norm, gn, gd;
   const double cf_tol = g_orc_cf_tol;
   double r_norm_0, cf_ave_0 = 0.0, cf_ave_1 = 0.0;
   int iter = 0, k, ret;
   if (converged) { *converged = 0; }
   orc_csr_matvec(n, L0->Ai, L0->Aj, L0->Aa, -1.0, x, 1.0, b, r0);
   memcpy(r, r0, sizeof(double) * (size_t) n);
   memcpy(p, r0, sizeof(double) * (size_t) n);
   b_norm = sqrt(orc_dot(n, b, b));
   res = orc_dot(n, r0, r0);
   r_norm = sqrt(res);
   r_norm_0 = r_norm;
   if (norms) { norms[0] = r_norm; }
   den_norm = b_norm > 0.0 ? b_norm : r_norm;
   epsilon = a_tol > tol * den_norm ? a_tol : tol * den_norm;
   if (g_orc_stop_crit) { epsilon = a_tol == 0.0 ? tol : a_tol; }     /* bicgstab.c:372-383 */
   if (rel_resid && b_norm > 0.0) { *rel_resid = r_norm / b_norm; }
   ret = 0;
   if (r_norm == 0.0) { goto done; }
   if (r_norm <= epsilon && iter >= min_iter) { if (converged) { *converged = 1; } goto done; }
   while (iter < max_iter)
   {
      iter++;
      orc_precond(h, precond_kind, p, v);
      orc_csr_matvec(n, L0->Ai, L0->Aj, L0->Aa, 1.0, v, 0.0, NULL, q);
      temp = orc_dot(n, r0, q);
      if (fabs(temp) >= epsmac) { alpha = res / temp; }
      else { ret = -iter - 1; goto done; }
      for (k = 0; k < n; k++) { x[k] += alpha * v[k]; }
      for (k = 0; k < n; k++) { r[k] += -alpha * q[k]; }
      orc_precond(h, precond_kind, r, v);
      orc_csr_matvec(n, L0->Ai, L0->Aj, L0->Aa, 1.0, v, 0.0, NULL, s);
      gn = orc_dot(n, r, s);
      gd = orc_dot(n, s, s);
      gamma = (gn == 0.0 && gd == 0.0) ? 0.0 : gn / gd;
      for (k = 0; k < n; k++) { x[k] += gamma * v[k]; }
      for (k = 0; k < n; k++) { r[k] += -gamma * s[k]; }
      r_norm = sqrt(orc_dot(n, r, r));
      if (norms) { norms[iter] = r_norm; }
      if (r_norm <= epsilon && iter >= min_iter)
      {
         orc_csr_matvec(n, L0->Ai, L0->Aj, L0->Aa, -1.0, x, 1.0, b, r);
         r_norm = sqrt(orc_dot(n, r, r));
         if (r_norm <= epsilon) { if (converged) { *converged = 1; } break; }
      }
      if (cf_tol > 0.0 && orc_cf_exceeded(cf_tol, r_norm, r_norm_0, iter, &cf_ave_0, &cf_ave_1)) { break; }
      if (fabs(res) >= epsmac) { beta = 1.0 / res; }
      else { ret = -iter - 1; goto done; }
      res = orc_dot(n, r0, r);
      beta *= res;
      for (k = 0; k < n; k++) { p[k] += -gamma * q[k]; }
      if (fabs(gamma) >= epsmac) { const double f = beta * alpha / gamma; for (k = 0; k < n; k++) { p[k] *= f; } }
      else { ret = -iter - 1; goto done; }
      for (k = 0; k < n; k++) { p[k] += r[k]; }
   }
   if (rel_resid) { *rel_resid = b_norm > 0.0 ? r_norm / b_norm : r_norm; }
   ret = iter;
done:
   free(r); free(r0); free(s); free(v); free(p); free(q);
   return ret;
}

/* hypre_LGMRESSolve (krylov/lgmres.c:337-...; ij -solver 51 AMG / 50 diagonal scaling) with cf_tol = 0:
 * restarted GMRES whose Krylov space is augmented with the last aug_dim error approximations (aug_vecs) and
 * their images (a_aug_vecs); approx_constant keeps the space at k_dim vectors.  Modified Gram-Schmidt, Givens
 * rotations and the true-residual check as in hypre_FlexGMRESSolve. */
int orc_lgmres(orc_amg *h, int precond_kind, const double *b, double *x, double tol, double a_tol, int k_dim,
               int aug_dim, int approx_constant, int min_iter, int max_iter, double *norms, double *rel_resid,
               int *converged)
{
   orc_level *L0 = &h->L[0];
   const int n = L0->n, kd = k_dim + aug_dim;
   const double epsmac = 1.e-16;
   const size_t nb = sizeof(double) * (size_t) n;
   double **p = (double **) calloc((size_t) k_dim + 1, sizeof(double *));
   double **av = (double **) calloc((size_t) aug_dim + 1, sizeof(double *));
   double **aav = (double **) calloc((size_t) aug_dim + 1, sizeof(double *));
   int *aug_order = (int *) calloc((size_t) aug_dim + 1, sizeof(int));
   double *r = (double *) calloc((size_t) n, sizeof(double));
   double *w = (double *) calloc((size_t) n, sizeof(double));
   double *rs = (double *) calloc((size_t) kd + 1, sizeof(double));
   double *c = (double *) calloc((size_t) kd + 1, sizeof(double));
   double *s = (double *) calloc((size_t) kd + 1, sizeof(double));
   double **hh = (double **) calloc((size_t) kd + 1, sizeof(double *));
   double b_norm, r_norm, r_norm_last = 0.0, den_norm, epsilon, t, gamma, tmp_norm;
   const double cf_tol = g_orc_cf_tol;
   double r_norm_0, cf_ave_0 = 0.0, cf_ave_1 = 0.0;
   int i, j, k, q, ii, iter = 0, aug_ct = 0, it_arnoldi, it_total, it_aug, order, spot = 0, break_value = 0;
   for (i = 0; i <= k_dim; i++) { p[i] = (double *) calloc((size_t) n, sizeof(double)); }
   for (i = 0; i <= aug_dim; i++) { av[i] = (double *) calloc((size_t) n, sizeof(double)); aav[i] = (double *) calloc((size_t) n, sizeof(double)); }
   for (i = 0; i <= kd; i++) { hh[i] = (double *) calloc((size_t) kd + 1, sizeof(double)); }
   if (converged) { *converged = 0; }
   if (rel_resid) { *rel_resid = 0.0; }
   orc_csr_matvec(n, L0->Ai, L0->Aj, L0->Aa, -1.0, x, 1.0, b, p[0]);
   b_norm = sqrt(orc_dot(n, b, b));
   r_norm = sqrt(orc_dot(n, p[0], p[0]));
   r_norm_0 = r_norm;
   if (norms) { norms[0] = r_norm; }
   den_norm = b_norm > 0.0 ? b_norm : r_norm;
   epsilon = a_tol > tol * den_norm ? a_tol : tol * den_norm;
   while (iter < max_iter)
   {
      rs[0] = r_norm;
      if (r_norm == 0.0) { goto done; }
      if (r_norm <= epsilon && iter >= min_iter)
      {
         orc_csr_matvec(n, L0->Ai, L0->Aj, L0->Aa, -1.0, x, 1.0, b, r);
         r_norm = sqrt(orc_dot(n, r, r));
         if (r_norm <= epsilon) { break; }
      }
      t = 1.0 / r_norm;
      r_norm_last = r_norm;
      for (q = 0; q < n; q++) { p[0][q] *= t; }
      i = 0;
      it_arnoldi = approx_constant ? k_dim - aug_ct : k_dim - aug_dim;
      it_total = it_arnoldi + aug_ct;
      it_aug = 0;
      while (i < it_total && iter < max_iter)
      {
         i++; iter++;
         if (i <= it_arnoldi)
         {
            orc_precond(h, precond_kind, p[i - 1], r);
            orc_csr_matvec(n, L0->Ai, L0->Aj, L0->Aa, 1.0, r, 0.0, NULL, p[i]);
         }
         else
         {
            it_aug++;
            order = i - it_arnoldi - 1;
            for (ii = 0; ii < aug_dim; ii++) { if (aug_order[ii] == order) { spot = ii; break; } }
            memcpy(p[i], aav[spot], nb);
         }
         for (j = 0; j < i; j++)
         {
            hh[j][i - 1] = orc_dot(n, p[j], p[i]);
            for (q = 0; q < n; q++) { p[i][q] += -hh[j][i - 1] * p[j][q]; }
         }
         t = sqrt(orc_dot(n, p[i], p[i]));
         hh[i][i - 1] = t;
         if (t != 0.0) { t = 1.0 / t; for (q = 0; q < n; q++) { p[i][q] *= t; } }
         for (j = 1; j < i; j++)
         {
            t = hh[j - 1][i - 1];
            hh[j - 1][i - 1] = s[j - 1] * hh[j][i - 1] + c[j - 1] * t;
            hh[j][i - 1] = -s[j - 1] * t + c[j - 1] * hh[j][i - 1];
         }
         t = hh[i][i - 1] * hh[i][i - 1];
         t += hh[i - 1][i - 1] * hh[i - 1][i - 1];
         gamma = sqrt(t);
         if (gamma == 0.0) { gamma = epsmac; }
         c[i - 1] = hh[i - 1][i - 1] / gamma;
         s[i - 1] = hh[i][i - 1] / gamma;
         rs[i] = -hh[i][i - 1] * rs[i - 1];
         rs[i] /= gamma;
         rs[i - 1] = c[i - 1] * rs[i - 1];
         hh[i - 1][i - 1] = s[i - 1] * hh[i][i - 1] + c[i - 1] * hh[i - 1][i - 1];
         r_norm = fabs(rs[i]);
         if (norms) { norms[iter] = r_norm; }
         if (cf_tol > 0.0 && orc_cf_exceeded(cf_tol, r_norm, r_norm_0, iter, &cf_ave_0, &cf_ave_1)) { break_value = 1; break; }
         if (r_norm <= epsilon && iter >= min_iter) { break; }
      }
      if (break_value) { break; }
      rs[i - 1] = rs[i - 1] / hh[i - 1][i - 1];
      for (k = i - 2; k >= 0; k--)
      {
         t = 0.0;
         for (j = k + 1; j < i; j++) { t -= hh[k][j] * rs[j]; }
         t += rs[k];
         rs[k] = t / hh[k][k];
      }
      if (it_arnoldi > i) { it_arnoldi = i; }
      if (!it_aug)
      {
         memcpy(w, p[i - 1], nb);
         for (q = 0; q < n; q++) { w[q] *= rs[i - 1]; }
         for (j = i - 2; j >= 0; j--) { for (q = 0; q < n; q++) { w[q] += rs[j] * p[j][q]; } }
      }
      else
      {
         memcpy(w, p[0], nb);
         for (q = 0; q < n; q++) { w[q] *= rs[0]; }
         for (j = 1; j < it_arnoldi; j++) { for (q = 0; q < n; q++) { w[q] += rs[j] * p[j][q]; } }
         for (ii = 0; ii < it_aug; ii++)
         {
            for (j = 0; j < aug_dim; j++) { if (aug_order[j] == ii) { spot = j; break; } }
            for (q = 0; q < n; q++) { w[q] += rs[it_arnoldi + ii] * av[spot][q]; }
         }
      }
      memcpy(av[aug_dim], w, nb);                 /* the new error approximation, before the preconditioner */
      orc_precond(h, precond_kind, w, r);
      for (q = 0; q < n; q++) { x[q] += r[q]; }
      if (r_norm <= epsilon && iter >= min_iter)
      {
         orc_csr_matvec(n, L0->Ai, L0->Aj, L0->Aa, -1.0, x, 1.0, b, r);
         r_norm = sqrt(orc_dot(n, r, r));
         if (r_norm <= epsilon) { if (converged) { *converged = 1; } break; }
         memcpy(p[0], r, nb);
         i = 0;
      }
      /* r0 (not scaled) into w, residual vector of the restart into p[0] */
      memcpy(w, p[0], nb);
      for (q = 0; q < n; q++) { w[q] *= r_norm_last; }
      for (j = i; j > 0; j--)
      {
         rs[j - 1] = -s[j - 1] * rs[j];
         rs[j] = c[j - 1] * rs[j];
      }
      if (i) { for (q = 0; q < n; q++) { p[i][q] += (rs[i] - 1.0) * p[i][q]; } }
      for (j = i - 1; j > 0; j--) { for (q = 0; q < n; q++) { p[i][q] += rs[j] * p[j][q]; } }
      if (i)
      {
         for (q = 0; q < n; q++) { p[0][q] += (rs[0] - 1.0) * p[0][q]; }
         for (q = 0; q < n; q++) { p[0][q] += p[i][q]; }
      }
      if (aug_dim > 0)
      {
         if (!aug_ct) { spot = 0; aug_ct++; }
         else if (aug_ct < aug_dim) { spot = aug_ct; aug_ct++; }
         else { for (ii = 0; ii < aug_dim; ii++) { if (aug_order[ii] == aug_dim - 1) { spot = ii; } } }
         memcpy(av[spot], av[aug_dim], nb);
         tmp_norm = 1.0 / sqrt(orc_dot(n, av[spot], av[spot]));
         for (q = 0; q < n; q++) { av[spot][q] *= tmp_norm; }
         for (ii = 0; ii < aug_dim; ii++) { aug_order[ii]++; }
         aug_order[spot] = 0;
         /* A * augvec = (r0 - rm) / norm : r0 is in w, rm in p[0] */
         memcpy(aav[spot], w, nb);
         for (q = 0; q < n; q++) { aav[spot][q] *= -1.0; }
         for (q = 0; q < n; q++) { aav[spot][q] += p[0][q]; }
         for (q = 0; q < n; q++) { aav[spot][q] *= -tmp_norm; }
      }
   }
   if (rel_resid) { *rel_resid = b_norm > 0.0 ? r_norm / b_norm : r_norm; }
done:
   for (i = 0; i <= k_dim; i++) { free(p[i]); }
   for (i = 0; i <= aug_dim; i++) { free(av[i]); free(aav[i]); }
   for (i = 0; i <= kd; i++) { free(hh[i]); }
   free(p); free(av); free(aav); free(aug_order); free(hh); free(r); free(w); free(rs); free(c); free(s);
   return iter;
}
