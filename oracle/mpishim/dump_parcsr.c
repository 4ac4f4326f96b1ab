/* oracle/mpishim/dump_parcsr.c -- TEST INFRASTRUCTURE.
 *
 * Runs the reference on N ranks (oracle/mpishim) and writes what every rank holds of the operators of a
 * BoomerAMG hierarchy: the ParCSR pieces (diag, offd, col_map_offd, row/column starts) and the matvec
 * communication package the reference negotiated with its real exchange
 * (hypre_MatvecCommPkgCreate, src/parcsr_mv/par_csr_communication.c:1163; assumed-partition
 * algorithm in new_commpkg.c).  tests/test_dist_host.py rebuilds the global matrices from these files
 * and checks that hypre_b200's communication-free restatement (hypre_b200/csrc/partition.cu,
 * b200_rowpart_create) produces, rank by rank, exactly the same blocks and the same plan.
 *
 *    mpirun_shim -np N dump_parcsr <7|27> nx ny nz P Q R outdir
 *
 * File <outdir>/<name>.<rank>.bin: int32 header {global_rows, global_cols, first_row, n_rows, first_col,
 * n_cols_diag, n_cols_offd, nnz_diag, nnz_offd, num_sends, n_send_elmts, num_recvs}, then int32 arrays
 * diag_i, diag_j, offd_i, offd_j, col_map_offd, send_procs, send_map_starts, send_map_elmts,
 * recv_procs, recv_vec_starts, then float64 arrays diag_a, offd_a.  Names: A0, P0, A1, P1, ...
 */
#include "_hypre_parcsr_ls.h"
#include <stdio.h>
#include <stdlib.h>

static void put_i(FILE *f, const HYPRE_Int *a, HYPRE_Int n) { if (n > 0) { fwrite(a, sizeof(HYPRE_Int), (size_t)n, f); } }

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
   /* the defaults src/test/ij.c passes (ij.c:248-302): HMIS, ext+i, Pmax 4 */
   HYPRE_Solver amg;
   HYPRE_BoomerAMGCreate(&amg);
   HYPRE_BoomerAMGSetCoarsenType(amg, 10);
   HYPRE_BoomerAMGSetInterpType(amg, 6);
   HYPRE_BoomerAMGSetPMaxElmts(amg, 4);
   HYPRE_BoomerAMGSetRelaxOrder(amg, 0);
   HYPRE_BoomerAMGSetRelaxType(amg, 18);
   HYPRE_BoomerAMGSetPrintLevel(amg, 0);
   HYPRE_BoomerAMGSetup(amg, A, b, x);
   hypre_ParAMGData *d = (hypre_ParAMGData *)amg;
   int nl = hypre_ParAMGDataNumLevels(d);
   for (int l = 0; l < nl; l++)
   {
      char name[32];
      snprintf(name, sizeof(name), "A%d", l);
      dump(hypre_ParAMGDataAArray(d)[l], dir, name, rank);
      if (l < nl - 1) { snprintf(name, sizeof(name), "P%d", l); dump(hypre_ParAMGDataPArray(d)[l], dir, name, rank); }
   }
   if (!rank) { printf("DUMP_OK levels %d ranks %d\n", nl, size); }
   HYPRE_BoomerAMGDestroy(amg);
   HYPRE_ParVectorDestroy(b); HYPRE_ParVectorDestroy(x);
   HYPRE_ParCSRMatrixDestroy(A);
   HYPRE_Finalize();
   MPI_Finalize();
   return 0;
}
