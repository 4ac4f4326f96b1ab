/* hypre_bridge.c -- plain-C accessors over hypre's host objects for the Python
 * side (ctypes cannot see hypre's structs).  No computation happens here: it
 * only hands out pointers / sizes of the reference's own arrays so that tests
 * and bench.py can (a) drive the HYPRE_* API the way src/test/ij.c does and
 * (b) feed the SAME hierarchy to the device path, the reference and the
 * plain-C oracle.  Compiled against the reference's headers.
 */
#include "_hypre_parcsr_ls.h"
#include "_hypre_parcsr_mv.h"
#include "_hypre_utilities.h"
#include "krylov.h"

/* b = value * ones with A's row partitioning (ij.c:3465-3510 builds rhs = ones, x0 = 0) */
void *hb_vector_create(void *A, double value)
{
   hypre_ParCSRMatrix *pA = (hypre_ParCSRMatrix *) A;
   hypre_ParVector *v = hypre_ParVectorCreate(hypre_ParCSRMatrixComm(pA),
                                              hypre_ParCSRMatrixGlobalNumRows(pA),
                                              hypre_ParCSRMatrixRowStarts(pA));
   hypre_ParVectorInitialize(v);
   hypre_ParVectorSetConstantValues(v, value);
   return v;
}
void *hb_vector_create_cols(void *A, double value)
{
   hypre_ParCSRMatrix *pA = (hypre_ParCSRMatrix *) A;
   hypre_ParVector *v = hypre_ParVectorCreate(hypre_ParCSRMatrixComm(pA),
                                              hypre_ParCSRMatrixGlobalNumCols(pA),
                                              hypre_ParCSRMatrixColStarts(pA));
   hypre_ParVectorInitialize(v);
   hypre_ParVectorSetConstantValues(v, value);
   return v;
}
void    hb_vector_destroy(void *v) { hypre_ParVectorDestroy((hypre_ParVector *) v); }
double *hb_vector_data(void *v) { return hypre_VectorData(hypre_ParVectorLocalVector((hypre_ParVector *) v)); }
int     hb_vector_size(void *v) { return hypre_VectorSize(hypre_ParVectorLocalVector((hypre_ParVector *) v)); }
void    hb_vector_set_all_zeros(void *v, int flag) { hypre_ParVectorAllZeros((hypre_ParVector *) v) = flag; }

void hb_matrix_blocks(void *A, int *n_rows, int *n_cols_diag, int **di, int **dj, double **da,
                      int *n_cols_offd, int **oi, int **oj, double **oa)
{
   hypre_ParCSRMatrix *pA = (hypre_ParCSRMatrix *) A;
   hypre_CSRMatrix *d = hypre_ParCSRMatrixDiag(pA), *o = hypre_ParCSRMatrixOffd(pA);
   *n_rows = hypre_CSRMatrixNumRows(d); *n_cols_diag = hypre_CSRMatrixNumCols(d);
   *di = hypre_CSRMatrixI(d); *dj = hypre_CSRMatrixJ(d); *da = hypre_CSRMatrixData(d);
   *n_cols_offd = hypre_CSRMatrixNumCols(o);
   *oi = hypre_CSRMatrixI(o); *oj = hypre_CSRMatrixJ(o); *oa = hypre_CSRMatrixData(o);
}
long long hb_matrix_global_rows(void *A) { return (long long) hypre_ParCSRMatrixGlobalNumRows((hypre_ParCSRMatrix *) A); }

/* ---- hypre_ParAMGData after HYPRE_BoomerAMGSetup */
int   hb_amg_num_levels(void *amg) { return hypre_ParAMGDataNumLevels((hypre_ParAMGData *) amg); }
void *hb_amg_A(void *amg, int l) { return hypre_ParAMGDataAArray((hypre_ParAMGData *) amg)[l]; }
void *hb_amg_P(void *amg, int l) { return hypre_ParAMGDataPArray((hypre_ParAMGData *) amg)[l]; }
void *hb_amg_F(void *amg, int l) { return hypre_ParAMGDataFArray((hypre_ParAMGData *) amg)[l]; }
void *hb_amg_U(void *amg, int l) { return hypre_ParAMGDataUArray((hypre_ParAMGData *) amg)[l]; }
void *hb_amg_Vtemp(void *amg) { return hypre_ParAMGDataVtemp((hypre_ParAMGData *) amg); }
void *hb_amg_Ztemp(void *amg) { return hypre_ParAMGDataZtemp((hypre_ParAMGData *) amg); }
void *hb_amg_Ptemp(void *amg) { return hypre_ParAMGDataPtemp((hypre_ParAMGData *) amg); }
void *hb_amg_Rtemp(void *amg) { return hypre_ParAMGDataRtemp((hypre_ParAMGData *) amg); }
double *hb_amg_l1(void *amg, int l)
{
   hypre_Vector **v = hypre_ParAMGDataL1Norms((hypre_ParAMGData *) amg);
   return (v && v[l]) ? hypre_VectorData(v[l]) : NULL;
}
double *hb_amg_cheby_ds(void *amg, int l)
{
   hypre_Vector **v = hypre_ParAMGDataChebyDS((hypre_ParAMGData *) amg);
   return (v && v[l]) ? hypre_VectorData(v[l]) : NULL;
}
double *hb_amg_cheby_coefs(void *amg, int l)
{
   HYPRE_Real **c = hypre_ParAMGDataChebyCoefs((hypre_ParAMGData *) amg);
   return (c && c[l]) ? c[l] : NULL;
}
/* CF_marker_array[l] of the setup (1 = C point, -1 = F point; NULL on the coarsest level) */
int *hb_amg_cf_marker(void *amg, int l)
{
   hypre_IntArray **m = hypre_ParAMGDataCFMarkerArray((hypre_ParAMGData *) amg);
   return (m && l < hypre_ParAMGDataNumLevels((hypre_ParAMGData *) amg) - 1 && m[l]) ? hypre_IntArrayData(m[l]) : NULL;
}
double hb_amg_relax_weight(void *amg, int l) { return hypre_ParAMGDataRelaxWeight((hypre_ParAMGData *) amg)[l]; }
double hb_amg_omega(void *amg, int l) { return hypre_ParAMGDataOmega((hypre_ParAMGData *) amg)[l]; }
/* out[0..3] grid_relax_type, out[4..7] num_grid_sweeps, out[8] cycle_type, out[9] fcycle,
 * out[10] cheby_order, out[11] cheby_scale, out[12] cheby_variant, out[13] relax_order */
void hb_amg_params(void *amg, int *out)
{
   hypre_ParAMGData *a = (hypre_ParAMGData *) amg;
   int k;
   for (k = 0; k < 4; k++)
   {
      out[k] = hypre_ParAMGDataGridRelaxType(a)[k];
      out[4 + k] = hypre_ParAMGDataNumGridSweeps(a)[k];
   }
   out[8] = hypre_ParAMGDataCycleType(a); out[9] = hypre_ParAMGDataFCycle(a);
   out[10] = hypre_ParAMGDataChebyOrder(a); out[11] = hypre_ParAMGDataChebyScale(a);
   out[12] = hypre_ParAMGDataChebyVariant(a); out[13] = hypre_ParAMGDataRelaxOrder(a);
}
/* dense coarse matrix of relax_type 9 (row major, N x N); builds it if the setup has not */
double *hb_amg_coarse_dense(void *amg, int *n)
{
   hypre_ParAMGData *a = (hypre_ParAMGData *) amg;
   int nl = hypre_ParAMGDataNumLevels(a);
   if (hypre_ParAMGDataGridRelaxType(a)[3] != 9 && hypre_ParAMGDataGridRelaxType(a)[3] != 19) { *n = 0; return NULL; }
   if (!hypre_ParAMGDataGSSetup(a)) { hypre_GaussElimSetup(a, nl - 1, hypre_ParAMGDataGridRelaxType(a)[3]); }
   *n = (int) hypre_ParCSRMatrixGlobalNumRows(hypre_ParAMGDataAArray(a)[nl - 1]);
   return hypre_ParAMGDataAMat(a);
}

/* ---- hypre_PCGData results / logging arrays (pcg.c:767-771,995-1003) */
double *hb_pcg_norms(void *pcg) { return ((hypre_PCGData *) pcg)->norms; }
double *hb_pcg_rel_norms(void *pcg) { return ((hypre_PCGData *) pcg)->rel_norms; }
void   *hb_pcg_precond_data(void *pcg) { return ((hypre_PCGData *) pcg)->precond_data; }

/* ---- hypre_GMRESData / hypre_BiCGSTABData logging arrays (gmres.c:610-630, bicgstab.c:470-474) */
double *hb_gmres_norms(void *g) { return ((hypre_GMRESData *) g)->norms; }
double *hb_flexgmres_norms(void *g) { return ((hypre_FlexGMRESData *) g)->norms; }
double *hb_lgmres_norms(void *g) { return ((hypre_LGMRESData *) g)->norms; }
double *hb_cogmres_norms(void *g) { return ((hypre_COGMRESData *) g)->norms; }
double *hb_bicgstab_norms(void *g) { return ((hypre_BiCGSTABData *) g)->norms; }
int     hb_bicgstab_converged(void *g) { return ((hypre_BiCGSTABData *) g)->converged; }

int  hb_error_flag(void) { return (int) hypre_error_flag; }
void hb_clear_errors(void) { HYPRE_ClearAllErrors(); }
