/* amg_oracle.c -- TEST INFRASTRUCTURE (never linked into the product).
 *
 * Plain-C, single-thread restatement of the reference's BoomerAMG-PCG solve phase
 * (hypre 2.32.0, paths under /root/reference/src).  Each function cites the code it
 * follows.  Parity pinned by tests/test_oracle.py against the compiled reference
 * (oracle/_ref/libHYPRE_ref.so) and the golden PCG histories in tests/golden/.
 */
#include "amg_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* y = alpha*A*x + beta*b ; row sums accumulated left to right in CSR storage order
 * (hypre_CSRMatrixMatvecOutOfPlaceHost, seq_mv/csr_matvec.c:683-847: y = scaled b, then
 * y += / -= / = alpha * tempx per row) */
void orc_csr_matvec(int n_rows, const int *ia, const int *ja, const double *a, double alpha,
                    const double *x, double beta, const double *b, double *y)
{
   int i, j;
   for (i = 0; i < n_rows; i++)
   {
      double tempx = 0.0;
      for (j = ia[i]; j < ia[i + 1]; j++) { tempx += a[j] * x[ja[j]]; }
      y[i] = (beta != 0.0 ? beta * b[i] : 0.0) + alpha * tempx;
   }
}

/* y = alpha*A^T*x + beta*y by scatter (hypre_CSRMatrixMatvecTHost, seq_mv/csr_matvec.c:913-1139:
 * y *= beta/alpha ; y[A_j[jj]] += A_data[jj]*x[i] ; y *= alpha) */
void orc_csr_matvecT(int n_rows, int n_cols, const int *ia, const int *ja, const double *a, double alpha,
                     const double *x, double beta, double *y)
{
   int i, j;
   double temp = beta / alpha;
   for (i = 0; i < n_cols; i++) { y[i] = (temp == 0.0) ? 0.0 : temp * y[i]; }
   for (i = 0; i < n_rows; i++)
   {
      for (j = ia[i]; j < ia[i + 1]; j++) { y[ja[j]] += a[j] * x[i]; }
   }
   if (alpha != 1.0) { for (i = 0; i < n_cols; i++) { y[i] *= alpha; } }
}

/* relax_type 18 / 7 (hypre_BoomerAMGRelax7Jacobi, parcsr_ls/par_relax.c:1177-1254):
 * Vtemp = f ; all_zeros ? Vtemp *= w : Vtemp = w*f - w*A*u ; u += Vtemp ./ l1 */
void orc_relax_l1jacobi(int n, const int *ia, const int *ja, const double *a, const double *f,
                        const double *l1, double w, double *u, double *vtemp, int all_zeros)
{
   int i;
   if (all_zeros) { for (i = 0; i < n; i++) { vtemp[i] = w * f[i]; } }
   else { orc_csr_matvec(n, ia, ja, a, -w, u, w, f, vtemp); }
   for (i = 0; i < n; i++) { u[i] += vtemp[i] / l1[i]; }
}

/* relax_type 16 (hypre_ParCSRRelax_Cheby_SolveHost, parcsr_ls/par_cheby.c:224-385) */
void orc_relax_cheby(int n, const int *ia, const int *ja, const double *a, const double *f,
                     const double *ds, const double *coefs, int order, int scale, double *u,
                     double *v, double *r, double *orig_u, double *tmp)
{
   int i, j, cheby_order;
   double mult;
   if (order > 4) { order = 4; }
   if (order < 1) { order = 1; }
   cheby_order = order - 1;
   if (!scale)
   {
      orc_csr_matvec(n, ia, ja, a, -1.0, u, 1.0, f, r);
      for (i = 0; i < n; i++) { orig_u[i] = u[i]; u[i] = r[i] * coefs[cheby_order]; }
      for (i = cheby_order - 1; i >= 0; i--)
      {
         orc_csr_matvec(n, ia, ja, a, 1.0, u, 0.0, NULL, v);
         mult = coefs[i];
         for (j = 0; j < n; j++) { u[j] = mult * r[j] + v[j]; }
      }
      for (i = 0; i < n; i++) { u[i] = orig_u[i] + u[i]; }
   }
   else
   {
      orc_csr_matvec(n, ia, ja, a, -1.0, u, 0.0, NULL, tmp);
      for (j = 0; j < n; j++) { r[j] = ds[j] * (f[j] + tmp[j]); }
      for (j = 0; j < n; j++) { orig_u[j] = u[j]; u[j] = r[j] * coefs[cheby_order]; }
      for (i = cheby_order - 1; i >= 0; i--)
      {
         for (j = 0; j < n; j++) { tmp[j] = ds[j] * u[j]; }
         orc_csr_matvec(n, ia, ja, a, 1.0, tmp, 0.0, NULL, v);
         mult = coefs[i];
         for (j = 0; j < n; j++) { u[j] = mult * r[j] + ds[j] * v[j]; }
      }
      for (j = 0; j < n; j++) { u[j] = orig_u[j] + ds[j] * u[j]; }
   }
}

/* relax_type 11 (inner 1) / 12 (inner 2)
 * (hypre_BoomerAMGRelaxTwoStageGaussSeidelHost, parcsr_ls/par_relax.c:1505-1588).
 * Returns 1 if a diagonal entry (first entry of a row) is zero (hypre_error_in_arg(1)). */
/* N-rank form: the inner stages use the strict lower part of each rank's DIAG block only (local
 * column < row, par_relax.c:1562-1581 loops over A_diag), i.e. entries whose column lies in the same
 * row block [starts[b], starts[b+1]) as the row.  nblocks <= 1 or starts == NULL: one block. */
int orc_relax_gs2_blocks(int n, const int *ia, const int *ja, const double *a, const double *f, double w,
                         int inner, double *u, double *vtemp, int nblocks, const int *starts)
{
   int i, k, jj, b = 0, err = 0;
   double multiplier = 1.0;
   if (nblocks <= 1 || !starts) { return orc_relax_gs2(n, ia, ja, a, f, w, inner, u, vtemp); }
   for (i = 0; i < n; i++) { if (a[ia[i]] == 0.0) { err = 1; } }
   orc_csr_matvec(n, ia, ja, a, -w, u, w, f, vtemp);
   for (i = 0; i < n; i++)
   {
      vtemp[i] /= a[ia[i]];
      u[i] += multiplier * vtemp[i];
   }
   multiplier *= -1.0;
   for (k = 0; k < inner; ++k)
   {
      b = nblocks - 1;
      for (i = n - 1; i >= 0; i--)
      {
         double res = 0.0;
         while (b > 0 && i < starts[b]) { b--; }
         for (jj = ia[i]; jj < ia[i + 1]; jj++)
         {
            int ii = ja[jj];
            if (ii < i && ii >= starts[b]) { res += a[jj] * vtemp[ii]; }
         }
         vtemp[i] = res / a[ia[i]];
         u[i] += multiplier * vtemp[i];
      }
      multiplier *= -1.0;
   }
   return err;
}

int orc_relax_gs2(int n, const int *ia, const int *ja, const double *a, const double *f, double w,
                  int inner, double *u, double *vtemp)
{
   int i, k, jj, err = 0;
   double multiplier = 1.0;
   for (i = 0; i < n; i++) { if (a[ia[i]] == 0.0) { err = 1; } }
   orc_csr_matvec(n, ia, ja, a, -w, u, w, f, vtemp);
   for (i = 0; i < n; i++)
   {
      vtemp[i] /= a[ia[i]];
      u[i] += multiplier * vtemp[i];
   }
   multiplier *= -1.0;
   for (k = 0; k < inner; ++k)
   {
      for (i = n - 1; i >= 0; i--)
      {
         double res = 0.0;
         for (jj = ia[i]; jj < ia[i + 1]; jj++)
         {
            int ii = ja[jj];
            if (ii < i) { res += a[jj] * vtemp[ii]; }
         }
         vtemp[i] = res / a[ia[i]];
         u[i] += multiplier * vtemp[i];
      }
      multiplier *= -1.0;
   }
   return err;
}

/* hypre_gselim (utilities/gselim.h:11-67): Gaussian elimination without pivoting,
 * A row-major n x n (overwritten), x = rhs in / solution out. */
int orc_gselim(double *A, double *x, int n)
{
   int j, k, m, error = 0;
   double factor, divA;
   if (n == 1)
   {
      if (A[0] != 0.0) { x[0] = x[0] / A[0]; } else { error++; }
      return error;
   }
   for (k = 0; k < n - 1; k++)
   {
      if (A[k * n + k] != 0.0)
      {
         divA = 1.0 / A[k * n + k];
         for (j = k + 1; j < n; j++)
         {
            if (A[j * n + k] != 0.0)
            {
               factor = A[j * n + k] * divA;
               for (m = k + 1; m < n; m++) { A[j * n + m] -= factor * A[k * n + m]; }
               x[j] -= factor * x[k];
            }
         }
      }
   }
   for (k = n - 1; k > 0; --k)
   {
      if (A[k * n + k] != 0.0)
      {
         x[k] /= A[k * n + k];
         for (j = 0; j < k; j++)
         {
            if (A[j * n + k] != 0.0) { x[j] -= x[k] * A[j * n + k]; }
         }
      }
   }
   if (A[0] != 0.0) { x[0] /= A[0]; }
   return error;
}
