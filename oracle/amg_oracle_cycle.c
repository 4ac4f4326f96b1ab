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
} orc_level;

struct orc_amg {
   int num_levels;
   orc_level *L;
   int relax_type[4], num_sweeps[4], cycle_type, fcycle, cheby_order, cheby_scale;
   int coarse_n;
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

/* one relaxation sweep on `level` (par_cycle.c:395-612 dispatch) */
static int orc_relax(orc_amg *h, int level, int relax_type, const double *f, double *u)
{
   orc_level *L = &h->L[level];
   int err = 0;
   switch (relax_type)
   {
      case 18: case 7:
         orc_relax_l1jacobi(L->n, L->Ai, L->Aj, L->Aa, f, L->l1, L->relax_weight, u, h->Vtemp, L->all_zeros);
         break;
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
         int e = orc_relax(h, level, relax_type, F, U);
         if (e < 0) { free(lev_counter); return e; }
         err |= e;
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
   else { orc_amg_cycle(h, b, p, 1); bi_prod = orc_dot(n, p, b); }
   eps = tol * tol;
   if (bi_prod <= 0.0) { memcpy(x, b, sizeof(double) * (size_t) n); free(p); free(s); free(r); return 0; }
   orc_csr_matvec(n, L0->Ai, L0->Aj, L0->Aa, -1.0, x, 1.0, b, r);
   memset(p, 0, sizeof(double) * (size_t) n);
   orc_amg_cycle(h, r, p, 1);
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
      memset(s, 0, sizeof(double) * (size_t) n);
      orc_amg_cycle(h, r, s, 1);
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
