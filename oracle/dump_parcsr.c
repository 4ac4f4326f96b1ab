/* oracle/dump_parcsr.c -- TEST INFRASTRUCTURE.
 *
 * Runs the reference on N ranks (hypre_b200/host/mpishim) and writes what every rank holds of the operators of a
 * BoomerAMG hierarchy: the ParCSR pieces (diag, offd, col_map_offd, row/column starts) and the matvec
 * communication package the reference negotiated with its real exchange
 * (hypre_MatvecCommPkgCreate, src/parcsr_mv/par_csr_communication.c:1163; assumed-partition
 * algorithm in new_commpkg.c).  tests/test_dist_host.py rebuilds the global matrices from these files
 * and checks that hypre_b200's communication-free restatement (hypre_b200/csrc/partition.cu,
 * b200_rowpart_create) produces, rank by rank, exactly the same blocks and the same plan.
 *
 *    mpirun_shim -np N dump_parcsr <7|27> nx ny nz P Q R outdir [relax_type [rhsrand [cycle_type [num_sweeps [agg_levels]]]]]
 *
 * With a relax type the program also does what `ij -solver 1 -rlx <type> [-rhsrand]` does -- BoomerAMG with
 * the driver's defaults (src/test/ij.c:248-302) as preconditioner of PCG, tol 1e-8 -- and writes what a
 * serial restatement needs to repeat that solve on the gathered hierarchy (tests/test_oracle_multirank.py):
 * per level and rank l1_<l>.<rank>.f64 / ds_<l>.<rank>.f64 (l1 norms, Chebyshev diagonal scaling), per level
 * coefs_<l>.f64, the gathered dense coarse matrix dense.f64, b.<rank>.f64, x.<rank>.f64 (solution),
 * norms.f64 / rel_norms.f64 (PCG log) and meta.txt.
 *
 * File <outdir>/<name>.<rank>.bin: int32 header {global_rows, global_cols, first_row, n_rows, first_col,
 * n_cols_diag, n_cols_offd, nnz_diag, nnz_offd, num_sends, n_send_elmts, num_recvs}, then int32 arrays
 * diag_i, diag_j, offd_i, offd_j, col_map_offd, send_procs, send_map_starts, send_map_elmts,
 * recv_procs, recv_vec_starts, then float64 arrays diag_a, offd_a.  Names: A0, P0, A1, P1, ...
 */
#include "_hypre_parcsr_ls.h"
#include <stdio.h>
#include <math.h>
#include <stdlib.h>

static void put_i(FILE *f, const HYPRE_Int *a, HYPRE_Int n) { if (n > 0) { fwrite(a, sizeof(HYPRE_Int), (size_t)n, f); } }

static void put_f64(const char *dir, const char *name, const HYPRE_Real *a, HYPRE_Int n)
{
   char path[1024];
   snprintf(path, sizeof(path), "%s/%s", dir, name);
   FILE *f = fopen(path, "wb");
   if (!f) { perror(path); MPI_Abort(MPI_COMM_WORLD, 4); }
   if (n > 0) { fwrite(a, sizeof(HYPRE_Real), (size_t)n, f); }
   fclose(f);
}

static void dump(hypre_ParCSRMatrix *A, const char *dir, const char *name, int rank)
{
   if (!hypre_ParCSRMatrixCommPkg(A)) { hypre_MatvecCommPkgCreate(A); }
   hypre_ParCSRCommPkg *pkg = hypre_ParCSRMatrixCommPkg(A);
   hypre_CSRMatrix *D = hypre_ParCSRMatrixDiag(A), *O = hypre_ParCSRMatrixOffd(A);
   HYPRE_Int n = hypre_CSRMatrixNumRows(D), nnzd = hypre_CSRMatrixI(D)[n], nco = hypre_CSRMatrixNumCols(O);
   HYPRE_Int nnzo = nco > 0 ? hypre_CSRMatrixI(O)[n] : 0;
   HYPRE_Int ns = hypre_ParCSRCommPkgNumSends(pkg), nr = hypre_ParCSRCommPkgNumRecvs(pkg);
   HYPRE_Int nse = ns > 0 ? hypre_ParCSRCommPkgSendMapStarts(pkg)[ns] : 0;
   HYPRE_Int hdr[12] = { (HYPRE_Int)hypre_ParCSRMatrixGlobalNumRows(A), (HYPRE_Int)hypre_ParCSRMatrixGlobalNumCols(A),
                         (HYPRE_Int)hypre_ParCSRMatrixFirstRowIndex(A), n, (HYPRE_Int)hypre_ParCSRMatrixFirstColDiag(A),
                         hypre_CSRMatrixNumCols(D), nco, nnzd, nnzo, ns, nse, nr };
   char path[1024];
   snprintf(path, sizeof(path), "%s/%s.%d.bin", dir, name, rank);
   FILE *f = fopen(path, "wb");
   if (!f) { perror(path); MPI_Abort(MPI_COMM_WORLD, 4); }
   fwrite(hdr, sizeof(HYPRE_Int), 12, f);
   put_i(f, hypre_CSRMatrixI(D), n + 1);
   put_i(f, hypre_CSRMatrixJ(D), nnzd);
   HYPRE_Int *zero = hypre_CTAlloc(HYPRE_Int, n + 1, HYPRE_MEMORY_HOST);
   put_i(f, nco > 0 ? hypre_CSRMatrixI(O) : zero, n + 1);
   put_i(f, hypre_CSRMatrixJ(O), nnzo);
   for (HYPRE_Int k = 0; k < nco; k++) { HYPRE_Int c = (HYPRE_Int)hypre_ParCSRMatrixColMapOffd(A)[k]; fwrite(&c, sizeof(HYPRE_Int), 1, f); }
   put_i(f, hypre_ParCSRCommPkgSendProcs(pkg), ns);
   if (ns > 0) { put_i(f, hypre_ParCSRCommPkgSendMapStarts(pkg), ns + 1); } else { put_i(f, zero, 1); }
   put_i(f, hypre_ParCSRCommPkgSendMapElmts(pkg), nse);
   put_i(f, hypre_ParCSRCommPkgRecvProcs(pkg), nr);
   if (nr > 0) { put_i(f, hypre_ParCSRCommPkgRecvVecStarts(pkg), nr + 1); } else { put_i(f, zero, 1); }
   if (nnzd > 0) { fwrite(hypre_CSRMatrixData(D), sizeof(HYPRE_Real), (size_t)nnzd, f); }
   if (nnzo > 0) { fwrite(hypre_CSRMatrixData(O), sizeof(HYPRE_Real), (size_t)nnzo, f); }
   hypre_TFree(zero, HYPRE_MEMORY_HOST);
   fclose(f);
}

int main(int argc, char **argv)
{
   int rank, size;
   MPI_Init(&argc, &argv);
   MPI_Comm_rank(MPI_COMM_WORLD, &rank);
   MPI_Comm_size(MPI_COMM_WORLD, &size);
   HYPRE_Initialize();
   if (argc < 9) { if (!rank) { fprintf(stderr, "usage: dump_parcsr <7|27> nx ny nz P Q R outdir\n"); } MPI_Abort(MPI_COMM_WORLD, 2); }
   int pts = atoi(argv[1]), nx = atoi(argv[2]), ny = atoi(argv[3]), nz = atoi(argv[4]);
   int P = atoi(argv[5]), Q = atoi(argv[6]), R = atoi(argv[7]);
   const char *dir = argv[8];
   int rlx = argc > 9 ? atoi(argv[9]) : -1, rhsrand = argc > 10 ? atoi(argv[10]) : 0;
   int cycle_type = argc > 11 ? atoi(argv[11]) : 1, sweeps = argc > 12 ? atoi(argv[12]) : 1, agg = argc > 13 ? atoi(argv[13]) : 0;
   if (P * Q * R != size) { if (!rank) { fprintf(stderr, "P*Q*R != ranks\n"); } MPI_Abort(MPI_COMM_WORLD, 2); }
   /* rank -> box coordinates as src/test/ij.c BuildParLaplacian does */
   int p = rank % P, q = ((rank - p) / P) % Q, r = (rank - p - P * q) / (P * Q);
   HYPRE_ParCSRMatrix A;
   if (pts == 27)
   {
      HYPRE_Real v[2] = { 26.0, -1.0 };
      if (nx == 1 || ny == 1 || nz == 1) { v[0] = 8.0; }
      if (nx * ny == 1 || nx * nz == 1 || ny * nz == 1) { v[0] = 2.0; }
      A = (HYPRE_ParCSRMatrix)GenerateLaplacian27pt(MPI_COMM_WORLD, nx, ny, nz, P, Q, R, p, q, r, v);
   }
   else
   {
      HYPRE_Real v[4] = { 0.0, -1.0, -1.0, -1.0 };
      if (nx > 1) { v[0] += 2.0; }
      if (ny > 1) { v[0] += 2.0; }
      if (nz > 1) { v[0] += 2.0; }
      A = (HYPRE_ParCSRMatrix)GenerateLaplacian(MPI_COMM_WORLD, nx, ny, nz, P, Q, R, p, q, r, v);
   }
   HYPRE_ParVector b, x;
   HYPRE_ParVectorCreate(MPI_COMM_WORLD, hypre_ParCSRMatrixGlobalNumRows((hypre_ParCSRMatrix *)A),
                         hypre_ParCSRMatrixRowStarts((hypre_ParCSRMatrix *)A), &b);
   HYPRE_ParVectorInitialize(b); HYPRE_ParVectorSetConstantValues(b, 1.0);
   HYPRE_ParVectorCreate(MPI_COMM_WORLD, hypre_ParCSRMatrixGlobalNumRows((hypre_ParCSRMatrix *)A),
                         hypre_ParCSRMatrixRowStarts((hypre_ParCSRMatrix *)A), &x);
   HYPRE_ParVectorInitialize(x); HYPRE_ParVectorSetConstantValues(x, 0.0);
   if (rhsrand)
   {
      /* ij -rhsrand (src/test/ij.c build_rhs_type 3): random values, scaled to unit 2-norm */
      HYPRE_Real nrm;
      HYPRE_ParVectorSetRandomValues(b, 22775);
      HYPRE_ParVectorInnerProd(b, b, &nrm);
      HYPRE_ParVectorScale(1.0 / sqrt(nrm), b);
   }
   /* the defaults src/test/ij.c passes (ij.c:248-302): HMIS, ext+i, Pmax 4, relax_order 0, max coarse 9 */
   HYPRE_Solver amg, pcg = NULL;
   HYPRE_BoomerAMGCreate(&amg);
   HYPRE_BoomerAMGSetCoarsenType(amg, 10);
   HYPRE_BoomerAMGSetInterpType(amg, 6);
   HYPRE_BoomerAMGSetPMaxElmts(amg, 4);
   HYPRE_BoomerAMGSetRelaxOrder(amg, 0);
   HYPRE_BoomerAMGSetMaxCoarseSize(amg, 9);
   HYPRE_BoomerAMGSetMinCoarseSize(amg, 0);
   HYPRE_BoomerAMGSetMaxRowSum(amg, 1.0);
   HYPRE_BoomerAMGSetStrongThreshold(amg, 0.25);
   HYPRE_BoomerAMGSetMaxLevels(amg, 25);
   HYPRE_BoomerAMGSetMeasureType(amg, 0);
   HYPRE_BoomerAMGSetNumSweeps(amg, sweeps);
   HYPRE_BoomerAMGSetCycleType(amg, cycle_type);
   HYPRE_BoomerAMGSetTruncFactor(amg, 0.0);
   HYPRE_BoomerAMGSetAggNumLevels(amg, agg);
   HYPRE_BoomerAMGSetPrintLevel(amg, 0);
   HYPRE_BoomerAMGSetChebyOrder(amg, 2);
   HYPRE_BoomerAMGSetChebyEigEst(amg, 10);
   HYPRE_BoomerAMGSetChebyVariant(amg, 0);
   HYPRE_BoomerAMGSetChebyScale(amg, 1);
   HYPRE_BoomerAMGSetChebyFraction(amg, 0.3);
   HYPRE_BoomerAMGSetRelaxType(amg, rlx >= 0 ? rlx : 18);
   int iters = 0;
   if (rlx >= 0)
   {
      HYPRE_BoomerAMGSetTol(amg, 0.0);
      HYPRE_BoomerAMGSetMaxIter(amg, 1);
      HYPRE_ParCSRPCGCreate(MPI_COMM_WORLD, &pcg);
      HYPRE_PCGSetMaxIter(pcg, 1000);
      HYPRE_PCGSetTol(pcg, 1.e-8);
      HYPRE_PCGSetTwoNorm(pcg, 1);
      HYPRE_PCGSetRelChange(pcg, 0);
      HYPRE_PCGSetPrintLevel(pcg, 0);
      HYPRE_PCGSetLogging(pcg, 1);
      HYPRE_PCGSetPrecond(pcg, (HYPRE_PtrToSolverFcn)HYPRE_BoomerAMGSolve, (HYPRE_PtrToSolverFcn)HYPRE_BoomerAMGSetup, amg);
      HYPRE_ParCSRPCGSetup(pcg, A, b, x);
      HYPRE_ParCSRPCGSolve(pcg, A, b, x);
      HYPRE_PCGGetNumIterations(pcg, &iters);
   }
   else { HYPRE_BoomerAMGSetup(amg, A, b, x); }
   hypre_ParAMGData *d = (hypre_ParAMGData *)amg;
   int nl = hypre_ParAMGDataNumLevels(d);
   for (int l = 0; l < nl; l++)
   {
      char name[64];
      snprintf(name, sizeof(name), "A%d", l);
      dump(hypre_ParAMGDataAArray(d)[l], dir, name, rank);
      if (l < nl - 1) { snprintf(name, sizeof(name), "P%d", l); dump(hypre_ParAMGDataPArray(d)[l], dir, name, rank); }
      if (rlx < 0) { continue; }
      HYPRE_Int n = hypre_CSRMatrixNumRows(hypre_ParCSRMatrixDiag(hypre_ParAMGDataAArray(d)[l]));
      hypre_Vector **l1 = hypre_ParAMGDataL1Norms(d), **ds = hypre_ParAMGDataChebyDS(d);
      HYPRE_Real **cf = hypre_ParAMGDataChebyCoefs(d);
      snprintf(name, sizeof(name), "l1_%d.%d.f64", l, rank);
      put_f64(dir, name, (l1 && l1[l]) ? hypre_VectorData(l1[l]) : NULL, (l1 && l1[l]) ? n : 0);
      snprintf(name, sizeof(name), "ds_%d.%d.f64", l, rank);
      put_f64(dir, name, (ds && ds[l]) ? hypre_VectorData(ds[l]) : NULL, (ds && ds[l]) ? n : 0);
      if (!rank)
      {
         snprintf(name, sizeof(name), "coefs_%d.f64", l);
         put_f64(dir, name, (cf && cf[l]) ? cf[l] : NULL, (cf && cf[l]) ? hypre_ParAMGDataChebyOrder(d) + 1 : 0);
      }
   }
   if (rlx >= 0)
   {
      char name[64];
      hypre_ParCSRMatrix *Ac = hypre_ParAMGDataAArray(d)[nl - 1];
      HYPRE_Int N = (HYPRE_Int)hypre_ParCSRMatrixGlobalNumRows(Ac);
      /* the gathered dense coarse matrix of relax type 9 lives on every rank that owns coarse rows */
      if (hypre_ParAMGDataGridRelaxType(d)[3] == 9 && hypre_ParAMGDataAMat(d) && hypre_ParCSRMatrixFirstRowIndex(Ac) == 0 &&
          hypre_CSRMatrixNumRows(hypre_ParCSRMatrixDiag(Ac)) > 0)
      {
         put_f64(dir, "dense.f64", hypre_ParAMGDataAMat(d), N * N);
      }
      HYPRE_Int nloc = hypre_CSRMatrixNumRows(hypre_ParCSRMatrixDiag((hypre_ParCSRMatrix *)A));
      snprintf(name, sizeof(name), "b.%d.f64", rank);
      put_f64(dir, name, hypre_VectorData(hypre_ParVectorLocalVector((hypre_ParVector *)b)), nloc);
      snprintf(name, sizeof(name), "x.%d.f64", rank);
      put_f64(dir, name, hypre_VectorData(hypre_ParVectorLocalVector((hypre_ParVector *)x)), nloc);
      if (!rank)
      {
         hypre_PCGData *pd = (hypre_PCGData *)pcg;
         put_f64(dir, "norms.f64", pd->norms, iters + 1);
         put_f64(dir, "rel_norms.f64", pd->rel_norms, iters + 1);
         char path[1024];
         snprintf(path, sizeof(path), "%s/meta.txt", dir);
         FILE *f = fopen(path, "w");
         fprintf(f, "levels %d\niterations %d\nfinal_rel %.15e\n", nl, iters, pd->rel_residual_norm);
         fprintf(f, "relax_type %d %d %d %d\n", hypre_ParAMGDataGridRelaxType(d)[0], hypre_ParAMGDataGridRelaxType(d)[1],
                 hypre_ParAMGDataGridRelaxType(d)[2], hypre_ParAMGDataGridRelaxType(d)[3]);
         fprintf(f, "num_sweeps %d %d %d %d\n", hypre_ParAMGDataNumGridSweeps(d)[0], hypre_ParAMGDataNumGridSweeps(d)[1],
                 hypre_ParAMGDataNumGridSweeps(d)[2], hypre_ParAMGDataNumGridSweeps(d)[3]);
         fprintf(f, "cycle_type %d\nfcycle %d\ncheby_order %d\ncheby_scale %d\n", hypre_ParAMGDataCycleType(d),
                 hypre_ParAMGDataFCycle(d), hypre_ParAMGDataChebyOrder(d), hypre_ParAMGDataChebyScale(d));
         fprintf(f, "relax_weight");
         for (int l = 0; l < nl; l++) { fprintf(f, " %.17g", hypre_ParAMGDataRelaxWeight(d)[l]); }
         fprintf(f, "\n");
         fclose(f);
      }
      HYPRE_ParCSRPCGDestroy(pcg);
   }
   if (!rank) { printf("DUMP_OK levels %d ranks %d\n", nl, size); }
   HYPRE_BoomerAMGDestroy(amg);
   HYPRE_ParVectorDestroy(b); HYPRE_ParVectorDestroy(x);
   HYPRE_ParCSRMatrixDestroy(A);
   HYPRE_Finalize();
   MPI_Finalize();
   return 0;
}
