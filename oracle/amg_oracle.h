/* amg_oracle.h -- TEST INFRASTRUCTURE: plain-C CPU restatement of the reference's
 * BoomerAMG-PCG solve phase (hypre 2.32.0).  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline leg may use it; it is never part of the product path.
 * Parity pinned: tests/test_oracle.py checks it against the compiled reference
 * (oracle/_ref/libHYPRE_ref.so) and against the golden PCG histories of BASELINE.md.
 */
#ifndef AMG_ORACLE_H
#define AMG_ORACLE_H
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_amg orc_amg;

void orc_csr_matvec(int n_rows, const int *ia, const int *ja, const double *a, double alpha,
                    const double *x, double beta, const double *b, double *y);
void orc_csr_matvecT(int n_rows, int n_cols, const int *ia, const int *ja, const double *a, double alpha,
                     const double *x, double beta, double *y);
void orc_relax_l1jacobi(int n, const int *ia, const int *ja, const double *a, const double *f,
                        const double *l1, double w, double *u, double *vtemp, int all_zeros);
void orc_relax_cheby(int n, const int *ia, const int *ja, const double *a, const double *f,
                     const double *ds, const double *coefs, int order, int scale, double *u,
                     double *v, double *r, double *orig_u, double *tmp);
int  orc_relax_gs2(int n, const int *ia, const int *ja, const double *a, const double *f, double w,
                   int inner, double *u, double *vtemp);
int  orc_relax_gs2_blocks(int n, const int *ia, const int *ja, const double *a, const double *f, double w,
                          int inner, double *u, double *vtemp, int nblocks, const int *starts);
int  orc_gselim(double *A, double *x, int n);

orc_amg *orc_amg_create(int num_levels);
void orc_amg_destroy(orc_amg *h);
/* arrays are borrowed (must outlive the object); P may be NULL on the last level */
void orc_amg_set_level(orc_amg *h, int level, int n, const int *Ai, const int *Aj, const double *Aa,
                       int n_coarse, const int *Pi, const int *Pj, const double *Pa,
                       const double *l1, const double *ds, const double *coefs, double relax_weight);
void orc_amg_set_cycle(orc_amg *h, const int relax_type[4], const int num_sweeps[4], int cycle_type,
                       int fcycle, int cheby_order, int cheby_scale);
void orc_amg_set_level_cf(orc_amg *h, int level, const int *cf_marker);   /* CF_marker_array[level], borrowed */
void orc_amg_set_relax_order(orc_amg *h, int relax_order);                  /* hypre_ParAMGDataRelaxOrder */
void orc_amg_set_coarse_dense(orc_amg *h, int n, const double *A_mat);
/* N-rank row partition of a level (nblocks+1 starts, copied): only two-stage GS depends on it */
void orc_amg_set_level_blocks(orc_amg *h, int level, int nblocks, const int *starts);
int  orc_amg_cycle(orc_amg *h, const double *f, double *u, int all_zeros);
/* cycles per preconditioner application (hypre_ParAMGDataMaxIter with tol 0; ij -precon_cycles), default 1 */
void orc_amg_set_precond_cycles(orc_amg *h, int cycles);
/* returns the iteration count; rel_norms[1..iters] = ||r||/||b|| (two_norm) */
int  orc_pcg(orc_amg *h, const double *b, double *x, double tol, int max_iter, int two_norm,
             double *norms, double *rel_norms, int *converged);

/* options read by the Krylov restatements below at the start of a solve: convergence-factor tolerance (all) and
 * BiCGSTAB's stop_crit; set, solve, reset to (0, 0) */
void orc_krylov_set_options(double cf_tol, int stop_crit);
/* precond_kind: 0 = one AMG cycle, 1 = diagonal scaling (HYPRE_ParCSRDiagScale), 2 = identity.
 * hypre_GMRESSolve (krylov/gmres.c:274-930, rel_change = 0, cf_tol = 0): returns the iteration count;
 * norms[0..iters] as the reference logs them. */
int  orc_gmres(orc_amg *h, int precond_kind, const double *b, double *x, double tol, double a_tol, int k_dim,
               int min_iter, int max_iter, int skip_real_r_check, double *norms, double *rel_resid,
               int *converged);
/* hypre_FlexGMRESSolve (krylov/flexgmres.c, default modify_pc): skip_real_r_check is ignored (the reference has none) */
int  orc_flexgmres(orc_amg *h, int precond_kind, const double *b, double *x, double tol, double a_tol, int k_dim,
                   int min_iter, int max_iter, int skip_real_r_check, double *norms, double *rel_resid,
                   int *converged);
/* hypre_LGMRESSolve (krylov/lgmres.c): GMRES augmented with aug_dim error approximations */
int  orc_lgmres(orc_amg *h, int precond_kind, const double *b, double *x, double tol, double a_tol, int k_dim,
                int aug_dim, int approx_constant, int min_iter, int max_iter, double *norms, double *rel_resid,
                int *converged);
/* hypre_COGMRESSolve (krylov/cogmres.c, cgs = 1): as orc_gmres with classical Gram-Schmidt */
int  orc_cogmres(orc_amg *h, int precond_kind, const double *b, double *x, double tol, double a_tol, int k_dim,
                 int min_iter, int max_iter, int skip_real_r_check, double *norms, double *rel_resid,
                 int *converged);
/* hypre_BiCGSTABSolve (krylov/bicgstab.c:226-640, stop_crit = 0, cf_tol = 0): iteration count, or
 * -(iter) - 1 where the reference reports a breakdown */
int  orc_bicgstab(orc_amg *h, int precond_kind, const double *b, double *x, double tol, double a_tol,
                  int min_iter, int max_iter, double *norms, double *rel_resid, int *converged);

#ifdef __cplusplus
}
#endif
#endif
