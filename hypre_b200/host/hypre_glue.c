/* hypre_glue.c -- the reference-side binding of the B200 solve phase.
 *
 * This file is what a hypre maintainer would add to the library: it reads
 * hypre's own host objects (hypre_ParCSRMatrix, hypre_ParAMGData, hypre_PCGData)
 * after the reference's host setup, uploads them once through the C ABI of
 * include/b200amg.h, and routes the solve-phase entry points to the device:
 *
 *   hypre_PCGSolve               (src/krylov/pcg.c:317)            -> b200_pcg_solve_host
 *   hypre_BoomerAMGSolve         (src/parcsr_ls/par_amg_solve.c:21) -> b200_amg_solve
 *   HYPRE_ParCSRMatrixMatvec     (src/parcsr_mv/HYPRE_parcsr_matrix.c) -> b200_parcsr_matvec_host
 *   hypre_BoomerAMGSetup/Destroy (mirror invalidation)
 *
 * It is built as libHYPRE_b200.so and placed BEFORE the reference library in
 * the link/load order, so these definitions interpose the reference's symbols
 * (ELF symbol interposition; the reference library is compiled -fPIC without
 * -Bsymbolic).  Calls that are not on the supported device path are rejected
 * loudly (HYPRE_ERROR_GENERIC) -- there is no CPU fallback for the solve --
 * except for Krylov solves whose operator table is not ParCSR (struct/sstruct
 * interfaces), which were never on this path and go to the reference untouched.
 *
 * Compiled against the reference's own headers (-I/root/reference/src/...);
 * no reference source is copied.
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

#include "_hypre_parcsr_ls.h"
#include "_hypre_parcsr_mv.h"
#include "_hypre_utilities.h"
#include "_hypre_IJ_mv.h"
#include "krylov.h"
#include "b200amg.h"

/* ------------------------------------------------------------------ state */
static b200_ctx *g_ctx = NULL;
static int g_verbose = -1;
/* test hook: 0 routes every interposed entry point to the reference (in-process A/B
 * comparison on the same hierarchy); the product default is 1 and nothing in the
 * product path ever clears it */
static int g_enabled = 1;
void hypre_b200_set_enabled(int enabled) { g_enabled = enabled; }
int  hypre_b200_get_enabled(void) { return g_enabled; }

/* distributed mode (one process per GPU, no MPI on the box): every process holds the SAME
 * single-rank hypre objects; process `g_rank` mirrors rows [first_row, first_row + n_local)
 * of every level on its GPU (hypre's 1-D row partition) and the solve runs over NCCL. */
static int g_rank = 0, g_nranks = 1;
/* native mode: this library was compiled against an MPI build of hypre (HYPRE_SEQUENTIAL undefined) and
 * the application runs on N ranks, one per GPU: hypre's objects ARE distributed -- every rank mirrors
 * what it owns, with the communication packages hypre negotiated (section 2 of INTEGRATION.md). */
static int g_native = 0;

typedef struct mirror_s {
   void            *key;      /* hypre_ParAMGData* or hypre_ParCSRMatrix* */
   b200_amg        *amg;
   b200_parcsr     *A;        /* level-0 operator (owned by amg when amg != NULL) */
   int              first_row, first_col;   /* this rank's offset into the global host vectors */
   int             *starts;   /* distributed mode: (levels) x (nranks+1) row starts; NULL otherwise */
   int              num_levels, repl_level;
   /* the host matrix the mirror was uploaded from (level 0 of a hierarchy) and its stamp at upload time: an
    * application that changes values in place (IJ SetValues / AddToValues + Assemble between time steps, or a
    * lagged preconditioner on a changed A) must not keep solving with the old device copy */
   hypre_ParCSRMatrix *src;
   unsigned long long  stamp;
   struct mirror_s *next;
} mirror_t;
static mirror_t *g_mirrors = NULL;

/* Cheap version stamp of a ParCSR matrix: array addresses, sizes and ~4096 sampled values of the diag / offd
 * blocks (FNV-1a).  Checked at every solve / SpMV entry (microseconds); an in-place update of the coefficients
 * changes it and the mirror is uploaded again.  An application that changes only entries the sample misses calls
 * hypre_b200_matrix_changed(A); HYPRE_IJMatrixAssemble does it by itself. */
static unsigned long long fnv_mix(unsigned long long h, const void *p, size_t n)
{
   const unsigned char *b = (const unsigned char *) p;
   size_t k;
   for (k = 0; k < n; k++) { h ^= b[k]; h *= 1099511628211ULL; }
   return h;
}
static unsigned long long matrix_stamp(hypre_ParCSRMatrix *A)
{
   unsigned long long h = 1469598103934665603ULL;
   int blk;
   for (blk = 0; blk < 2; blk++)
   {
      hypre_CSRMatrix *M = blk == 0 ? hypre_ParCSRMatrixDiag(A) : hypre_ParCSRMatrixOffd(A);
      const HYPRE_Real *a = hypre_CSRMatrixData(M);
      const HYPRE_Int  *ia = hypre_CSRMatrixI(M);
      HYPRE_Int n = hypre_CSRMatrixNumRows(M);
      HYPRE_Int nnz = (ia && n > 0) ? ia[n] : 0;
      HYPRE_Int step = nnz / 2048 + 1, k;
      h = fnv_mix(h, &a, sizeof(a)); h = fnv_mix(h, &ia, sizeof(ia)); h = fnv_mix(h, &nnz, sizeof(nnz));
      for (k = 0; a && k < nnz; k += step) { h = fnv_mix(h, a + k, sizeof(HYPRE_Real)); }
      if (a && nnz > 0) { h = fnv_mix(h, a + nnz - 1, sizeof(HYPRE_Real)); }
   }
   return h;
}

static int verbose(void)
{
   if (g_verbose < 0) { const char *e = getenv("HYPRE_B200_VERBOSE"); g_verbose = e ? atoi(e) : 0; }
   return g_verbose;
}

static b200_ctx *glue_ctx(void)
{
#ifndef HYPRE_SEQUENTIAL
   if (!g_ctx)
   {
      /* first use is collective (a solve): rank 0 creates the NCCL id, hypre's own MPI layer carries it */
      HYPRE_Int np = 1, id = 0;
      hypre_MPI_Comm_size(hypre_MPI_COMM_WORLD, &np);
      hypre_MPI_Comm_rank(hypre_MPI_COMM_WORLD, &id);
      if (np > 1)
      {
         char uid[128];
         const char *e = getenv("HYPRE_B200_DEVICE");
         int dev = e ? atoi(e) : (getenv("LOCAL_RANK") ? atoi(getenv("LOCAL_RANK")) : (int) id);
         memset(uid, 0, sizeof(uid));
         if (id == 0 && b200_nccl_unique_id(uid) != B200_OK)
         {
            fprintf(stderr, "hypre_b200: cannot create the NCCL id: %s\n", b200_last_error());
         }
         hypre_MPI_Bcast(uid, 128, hypre_MPI_BYTE, 0, hypre_MPI_COMM_WORLD);
         if (b200_ctx_create_dist(dev, (int) id, (int) np, uid, &g_ctx) != B200_OK)
         {
            fprintf(stderr, "hypre_b200: rank %d cannot create the distributed context: %s\n", (int) id, b200_last_error());
            g_ctx = NULL;
            return NULL;
         }
         g_rank = (int) id; g_nranks = (int) np; g_native = 1;
         return g_ctx;
      }
   }
#endif
   if (!g_ctx)
   {
      int dev = 0;
      const char *e = getenv("HYPRE_B200_DEVICE");
      if (!e) { e = getenv("LOCAL_RANK"); }
      if (e) { dev = atoi(e); }
      if (b200_ctx_create(dev, &g_ctx) != B200_OK)
      {
         fprintf(stderr, "hypre_b200: cannot create device context: %s\n", b200_last_error());
         g_ctx = NULL;
      }
   }
   return g_ctx;
}

b200_ctx *hypre_b200_context(void) { return glue_ctx(); }

/* Join an NCCL communicator of `nranks` processes (uid from b200_nccl_unique_id on rank 0,
 * broadcast by the caller).  Must be called before anything is mirrored. */
int hypre_b200_init_dist(int device, int rank, int nranks, const char *uid)
{
   if (g_ctx || g_mirrors) { fprintf(stderr, "hypre_b200_init_dist: context already exists\n"); return -1; }
   if (b200_ctx_create_dist(device, rank, nranks, uid, &g_ctx) != B200_OK)
   {
      fprintf(stderr, "hypre_b200: cannot create distributed context: %s\n", b200_last_error());
      g_ctx = NULL;
      return -1;
   }
   g_rank = rank; g_nranks = nranks;
   return 0;
}
int hypre_b200_rank(void) { return g_rank; }
int hypre_b200_nranks(void) { return g_nranks; }

static mirror_t *mirror_find(void *key)
{
   mirror_t *m;
   for (m = g_mirrors; m; m = m->next) { if (m->key == key) { return m; } }
   return NULL;
}

static void mirror_drop(void *key)
{
   mirror_t **pp = &g_mirrors;
   while (*pp)
   {
      if ((*pp)->key == key)
      {
         mirror_t *m = *pp;
         *pp = m->next;
         if (m->amg) { b200_amg_destroy(m->amg); }
         else if (m->A) { b200_parcsr_destroy(m->A); }
         free(m->starts);
         free(m);
         return;
      }
      pp = &(*pp)->next;
   }
}

void hypre_b200_drop_vector_bindings(void);
void hypre_b200_reset(void)
{
   while (g_mirrors) { mirror_drop(g_mirrors->key); }
   hypre_b200_drop_vector_bindings();
   if (g_ctx) { b200_ctx_destroy(g_ctx); g_ctx = NULL; }
}

#define GLUE_FAIL(msg)                                                              \
   do {                                                                             \
      char _b[1200];                                                                \
      snprintf(_b, sizeof(_b), "hypre_b200: %s (%s)", msg, b200_last_error());      \
      fprintf(stderr, "%s\n", _b);                                                  \
      hypre_error_w_msg(HYPRE_ERROR_GENERIC, _b);                                   \
   } while (0)

/* ------------------------------------------------- device-resident vectors
 * SURVEY 8(f)-2.  The reference passes device vectors through the memory location of the hypre_Vector
 * (HYPRE_IJVectorInitialize_v2(.., HYPRE_MEMORY_DEVICE), ij -memory_device; docs/usr-manual/solvers-boomeramg.rst:
 * 217-253).  The host-only build of the reference this binding sits on has no device memory location, so an
 * application states it here instead: it BINDS a device array (this rank's local part, b200_dev_alloc or any
 * cudaMalloc pointer on the binding's device) to a hypre_ParVector.  The Krylov bindings below then solve on
 * those arrays in place -- no host<->device copy of b or x -- and leave the vector's host array untouched until
 * hypre_b200_vector_sync_to_host is called. */
typedef struct devvec_s { void *key; double *dev; struct devvec_s *next; } devvec_t;
static devvec_t *g_devvecs = NULL;

static double *devvec_find(void *key)
{
   devvec_t *d;
   for (d = g_devvecs; d; d = d->next) { if (d->key == key) { return d->dev; } }
   return NULL;
}

/* dev == NULL removes the binding.  Returns 0. */
int hypre_b200_vector_bind_device(void *par_vector, double *dev)
{
   devvec_t **pp = &g_devvecs, *d;
   while (*pp && (*pp)->key != par_vector) { pp = &(*pp)->next; }
   if (*pp)
   {
      if (dev) { (*pp)->dev = dev; }
      else { d = *pp; *pp = d->next; free(d); }
      return 0;
   }
   if (!dev) { return 0; }
   d = (devvec_t *) calloc(1, sizeof(devvec_t));
   d->key = par_vector; d->dev = dev; d->next = g_devvecs; g_devvecs = d;
   return 0;
}
double *hypre_b200_vector_device(void *par_vector) { return devvec_find(par_vector); }

/* copy entries [first, first + n) of the vector's host array from / to the bound device array (n < 0: the whole
 * local vector).  Returns 0, or -1 when the vector is not bound or the copy fails. */
static int devvec_sync(void *par_vector, int first, int n, int to_host)
{
   hypre_Vector *loc = hypre_ParVectorLocalVector((hypre_ParVector *) par_vector);
   double *dev = devvec_find(par_vector);
   b200_ctx *ctx = glue_ctx();
   if (!dev || !ctx) { return -1; }
   if (n < 0) { first = 0; n = (int) hypre_VectorSize(loc); }
   if (to_host) { return b200_d2h(ctx, hypre_VectorData(loc) + first, dev, n) == B200_OK ? 0 : -1; }
   return b200_h2d(ctx, dev, hypre_VectorData(loc) + first, n) == B200_OK ? 0 : -1;
}
int hypre_b200_vector_sync_to_host(void *par_vector, int first, int n) { return devvec_sync(par_vector, first, n, 1); }
int hypre_b200_vector_sync_to_device(void *par_vector, int first, int n) { return devvec_sync(par_vector, first, n, 0); }

void hypre_b200_drop_vector_bindings(void)
{
   while (g_devvecs) { devvec_t *d = g_devvecs; g_devvecs = d->next; free(d); }
}

/* operands of a Krylov binding: both vectors bound to device arrays (solve in place), or both on the host */
static int krylov_operands(hypre_ParVector *pb, hypre_ParVector *px, double **db, double **dx)
{
   *db = devvec_find(pb); *dx = devvec_find(px);
   if ((*db == NULL) != (*dx == NULL))
   {
      GLUE_FAIL("both vector operands must be bound to device arrays (hypre_b200_vector_bind_device) or both live on the host");
      return -1;
   }
   return 0;
}

/* --------------------------------------------------------------- uploads */
/* hypre_ParCSRMatrix -> b200_parcsr (diag, offd, halo plan of the CommPkg) */
static b200_parcsr *upload_parcsr(hypre_ParCSRMatrix *A)
{
   hypre_CSRMatrix     *diag = hypre_ParCSRMatrixDiag(A);
   hypre_CSRMatrix     *offd = hypre_ParCSRMatrixOffd(A);
   hypre_ParCSRCommPkg *pkg;
   b200_parcsr         *dA = NULL;
   HYPRE_Int            n_offd = hypre_CSRMatrixNumCols(offd);
   int ns = 0, nr = 0;
   const int *sp = NULL, *sms = NULL, *sme = NULL, *rp = NULL, *rvs = NULL;

   if (!hypre_ParCSRMatrixCommPkg(A)) { hypre_MatvecCommPkgCreate(A); }
   pkg = hypre_ParCSRMatrixCommPkg(A);
   if (pkg)
   {
      ns  = hypre_ParCSRCommPkgNumSends(pkg);
      nr  = hypre_ParCSRCommPkgNumRecvs(pkg);
      sp  = hypre_ParCSRCommPkgSendProcs(pkg);
      sms = hypre_ParCSRCommPkgSendMapStarts(pkg);
      sme = hypre_ParCSRCommPkgSendMapElmts(pkg);
      rp  = hypre_ParCSRCommPkgRecvProcs(pkg);
      rvs = hypre_ParCSRCommPkgRecvVecStarts(pkg);
   }
   if (b200_parcsr_create(glue_ctx(), hypre_CSRMatrixNumRows(diag), hypre_CSRMatrixNumCols(diag),
                          hypre_CSRMatrixI(diag), hypre_CSRMatrixJ(diag), hypre_CSRMatrixData(diag),
                          n_offd, hypre_CSRMatrixI(offd), hypre_CSRMatrixJ(offd), hypre_CSRMatrixData(offd),
                          ns, sp, sms, sme, nr, rp, rvs, &dA) != B200_OK)
   {
      return NULL;
   }
   return dA;
}

/* hypre_GeneratePartitioning (src/parcsr_mv/par_csr_matrix.c / utilities): n rows over P ranks,
 * the first n % P ranks get one extra row */
static void even_starts(int n, int P, int *starts)
{
   int size = n / P, rest = n - size * P, i;
   starts[0] = 0;
   for (i = 0; i < P; i++) { starts[i + 1] = starts[i] + size + (i < rest ? 1 : 0); }
}

/* rows [rs[g_rank], rs[g_rank+1]) of a single-rank hypre_ParCSRMatrix -> b200_parcsr with the
 * halo plan of that partition (b200_rowpart_create = GenerateDiagAndOffd + MatvecCommPkgCreate) */
static b200_parcsr *upload_parcsr_rows(hypre_ParCSRMatrix *A, const int *rs, const int *cs)
{
   hypre_CSRMatrix *diag = hypre_ParCSRMatrixDiag(A);
   b200_rowpart    *part = NULL;
   b200_parcsr     *dA = NULL;
   if (hypre_CSRMatrixNumCols(hypre_ParCSRMatrixOffd(A)) != 0)
   {
      fprintf(stderr, "hypre_b200: distributed mode expects single-rank (replicated) hypre matrices\n");
      return NULL;
   }
   if (b200_rowpart_create(g_rank, g_nranks, hypre_CSRMatrixNumRows(diag), hypre_CSRMatrixNumCols(diag),
                           hypre_CSRMatrixI(diag), hypre_CSRMatrixJ(diag), hypre_CSRMatrixData(diag),
                           rs, cs, &part) != B200_OK) { return NULL; }
   if (b200_parcsr_create_from_rowpart(glue_ctx(), part, &dA) != B200_OK) { dA = NULL; }
   b200_rowpart_destroy(part);
   return dA;
}

/* rows [r0, r1) of a single-rank hypre_ParCSRMatrix with ALL (global) columns, as a communication-free
 * single-block b200_parcsr: interpolation onto a replicated coarse level */
static b200_parcsr *upload_parcsr_row_slab(hypre_ParCSRMatrix *A, int r0, int r1)
{
   hypre_CSRMatrix *diag = hypre_ParCSRMatrixDiag(A);
   const HYPRE_Int *ai = hypre_CSRMatrixI(diag);
   b200_parcsr *dA = NULL;
   int n = r1 - r0, k;
   int *ri = (int *) malloc(sizeof(int) * (size_t)(n + 1));
   for (k = 0; k <= n; k++) { ri[k] = ai[r0 + k] - ai[r0]; }
   if (b200_parcsr_create(glue_ctx(), n, hypre_CSRMatrixNumCols(diag), ri, hypre_CSRMatrixJ(diag) + ai[r0],
                          hypre_CSRMatrixData(diag) + ai[r0], 0, NULL, NULL, NULL, 0, NULL, NULL, NULL, 0, NULL, NULL,
                          &dA) != B200_OK) { dA = NULL; }
   free(ri);
   return dA;
}

/* hypre_ParAMGData (after the reference's host setup) -> b200_amg */
/* ---- native mode, replicated coarse levels: a small level is gathered on every rank and runs without communication
 * (the reference's answer to idle ranks on coarse grids is seq_threshold / hypre_seqAMGCycle, par_cycle.c:362-369;
 * hypre_ParCSRMatrixToCSRMatrixAll leaves out the ranks that own no rows, so the gather is done here). */
#ifndef HYPRE_SEQUENTIAL
/* this rank's rows of A with GLOBAL column indices, diag entries first (the diagonal stays the first entry) */
static void local_rows_global_cols(hypre_ParCSRMatrix *A, int **ri, int **rj, double **ra)
{
   hypre_CSRMatrix *d = hypre_ParCSRMatrixDiag(A), *o = hypre_ParCSRMatrixOffd(A);
   const HYPRE_Int *di = hypre_CSRMatrixI(d), *dj = hypre_CSRMatrixJ(d), *oi = hypre_CSRMatrixI(o), *oj = hypre_CSRMatrixJ(o);
   const HYPRE_Real *da = hypre_CSRMatrixData(d), *oa = hypre_CSRMatrixData(o);
   const HYPRE_BigInt *cmap = hypre_ParCSRMatrixColMapOffd(A);
   const HYPRE_BigInt c0 = hypre_ParCSRMatrixFirstColDiag(A);
   const HYPRE_Int n = hypre_CSRMatrixNumRows(d), no = hypre_CSRMatrixNumCols(o);
   HYPRE_Int r, k, nnz = (n > 0 ? di[n] : 0) + ((n > 0 && no > 0) ? oi[n] : 0), p = 0;
   *ri = (int *) calloc((size_t) n + 1, sizeof(int));
   *rj = (int *) malloc(sizeof(int) * (size_t) (nnz > 0 ? nnz : 1));
   *ra = (double *) malloc(sizeof(double) * (size_t) (nnz > 0 ? nnz : 1));
   for (r = 0; r < n; r++)
   {
      for (k = di[r]; k < di[r + 1]; k++) { (*rj)[p] = (int) (c0 + dj[k]); (*ra)[p++] = da[k]; }
      if (no > 0) { for (k = oi[r]; k < oi[r + 1]; k++) { (*rj)[p] = (int) cmap[oj[k]]; (*ra)[p++] = oa[k]; } }
      (*ri)[r + 1] = p;
   }
}

/* the whole matrix on every rank: rows in global order; rows_of[r] = rows owned by rank r (returned for the vectors) */
static b200_parcsr *upload_parcsr_gathered(hypre_ParCSRMatrix *A, int *rows_of)
{
   MPI_Comm comm = hypre_ParCSRMatrixComm(A);
   int *li, *lj, *cnt, *dsp, *ncnt, *ndsp, *gi, *gj, *glen, r, n_loc, nnz_loc, n_all = 0, nnz_all = 0;
   double *la, *ga;
   b200_parcsr *dA = NULL;
   local_rows_global_cols(A, &li, &lj, &la);
   n_loc = (int) hypre_CSRMatrixNumRows(hypre_ParCSRMatrixDiag(A));
   nnz_loc = li[n_loc];
   cnt = (int *) calloc((size_t) g_nranks, sizeof(int)); dsp = (int *) calloc((size_t) g_nranks + 1, sizeof(int));
   ncnt = (int *) calloc((size_t) g_nranks, sizeof(int)); ndsp = (int *) calloc((size_t) g_nranks + 1, sizeof(int));
   hypre_MPI_Allgather(&n_loc, 1, HYPRE_MPI_INT, cnt, 1, HYPRE_MPI_INT, comm);
   hypre_MPI_Allgather(&nnz_loc, 1, HYPRE_MPI_INT, ncnt, 1, HYPRE_MPI_INT, comm);
   for (r = 0; r < g_nranks; r++) { dsp[r + 1] = dsp[r] + cnt[r]; ndsp[r + 1] = ndsp[r] + ncnt[r]; if (rows_of) { rows_of[r] = cnt[r]; } }
   n_all = dsp[g_nranks]; nnz_all = ndsp[g_nranks];
   glen = (int *) calloc((size_t) n_all + 1, sizeof(int));
   gi = (int *) calloc((size_t) n_all + 1, sizeof(int));
   gj = (int *) malloc(sizeof(int) * (size_t) (nnz_all > 0 ? nnz_all : 1));
   ga = (double *) malloc(sizeof(double) * (size_t) (nnz_all > 0 ? nnz_all : 1));
   {
      int *llen = (int *) calloc((size_t) n_loc + 1, sizeof(int));
      for (r = 0; r < n_loc; r++) { llen[r] = li[r + 1] - li[r]; }
      hypre_MPI_Allgatherv(llen, n_loc, HYPRE_MPI_INT, glen, cnt, dsp, HYPRE_MPI_INT, comm);
      free(llen);
   }
   hypre_MPI_Allgatherv(lj, nnz_loc, HYPRE_MPI_INT, gj, ncnt, ndsp, HYPRE_MPI_INT, comm);
   hypre_MPI_Allgatherv(la, nnz_loc, HYPRE_MPI_REAL, ga, ncnt, ndsp, HYPRE_MPI_REAL, comm);
   for (r = 0; r < n_all; r++) { gi[r + 1] = gi[r] + glen[r]; }
   if (b200_parcsr_create(glue_ctx(), n_all, (int) hypre_ParCSRMatrixGlobalNumCols(A), gi, gj, ga, 0, NULL, NULL, NULL,
                          0, NULL, NULL, NULL, 0, NULL, NULL, &dA) != B200_OK) { dA = NULL; }
   free(li); free(lj); free(la); free(cnt); free(dsp); free(ncnt); free(ndsp); free(glen); free(gi); free(gj); free(ga);
   return dA;
}

/* this rank's rows of P with global (replicated) coarse columns: restriction onto the first replicated level ends with
 * a sum over ranks, prolongation from it is local */
static b200_parcsr *upload_parcsr_local_slab(hypre_ParCSRMatrix *P)
{
   int *li, *lj;
   double *la;
   b200_parcsr *dP = NULL;
   local_rows_global_cols(P, &li, &lj, &la);
   if (b200_parcsr_create(glue_ctx(), (int) hypre_CSRMatrixNumRows(hypre_ParCSRMatrixDiag(P)),
                          (int) hypre_ParCSRMatrixGlobalNumCols(P), li, lj, la, 0, NULL, NULL, NULL,
                          0, NULL, NULL, NULL, 0, NULL, NULL, &dP) != B200_OK) { dP = NULL; }
   free(li); free(lj); free(la);
   return dP;
}

/* a level vector (l1 norms, Chebyshev scaling) of a replicated level: every rank's part, in rank order */
static double *gather_level_vector(MPI_Comm comm, const double *local, const int *rows_of)
{
   int r, *dsp = (int *) calloc((size_t) g_nranks + 1, sizeof(int));
   double *all, dummy = 0.0;
   for (r = 0; r < g_nranks; r++) { dsp[r + 1] = dsp[r] + rows_of[r]; }
   all = (double *) calloc((size_t) (dsp[g_nranks] > 0 ? dsp[g_nranks] : 1), sizeof(double));
   hypre_MPI_Allgatherv((void *) (local ? local : &dummy), rows_of[g_rank], HYPRE_MPI_REAL, all, (HYPRE_Int *) rows_of, dsp, HYPRE_MPI_REAL, comm);
   free(dsp);
   return all;
}
#endif

static b200_amg *upload_amg(hypre_ParAMGData *amg_data, int **starts_out, int *repl_out)
{
   b200_ctx  *ctx = glue_ctx();
   b200_amg  *amg = NULL;
   HYPRE_Int  nl = hypre_ParAMGDataNumLevels(amg_data);
   hypre_ParCSRMatrix **A_array = hypre_ParAMGDataAArray(amg_data);
   hypre_ParCSRMatrix **P_array = hypre_ParAMGDataPArray(amg_data);
   HYPRE_Int *grt = hypre_ParAMGDataGridRelaxType(amg_data);
   HYPRE_Int *ngs = hypre_ParAMGDataNumGridSweeps(amg_data);
   hypre_Vector **l1 = hypre_ParAMGDataL1Norms(amg_data);
   hypre_Vector **ds = hypre_ParAMGDataChebyDS(amg_data);
   HYPRE_Real **coefs = hypre_ParAMGDataChebyCoefs(amg_data);
   int relax_type[4], num_sweeps[4], l, k, uses_cheby = 0, uses_jac0 = 0;
   int repl = 1 << 30;   /* distributed mode: first level replicated on every rank */
   int *starts = NULL;   /* distributed mode: (nl+1) x (g_nranks+1) row starts per level */
   int *rows_of = NULL;  /* native mode: (nl) x (g_nranks) rows owned by every rank on the replicated levels */
#define STARTS(l) (starts + (size_t)(l) * (g_nranks + 1))

   if (!ctx) { return NULL; }
   /* options with no device implementation are rejected, never silently run on the CPU */
   if (hypre_ParAMGDataBlockMode(amg_data) || hypre_ParAMGDataSmoothNumLevels(amg_data) > 0 ||
       hypre_ParAMGDataAdditive(amg_data) >= 0 || hypre_ParAMGDataMultAdditive(amg_data) >= 0 ||
       hypre_ParAMGDataSimple(amg_data) >= 0 || hypre_ParAMGDataRestriction(amg_data) ||
       (hypre_ParAMGDataRelaxOrder(amg_data) != 0 && hypre_ParAMGDataRelaxOrder(amg_data) != 1) ||
       hypre_ParAMGDataGridRelaxPoints(amg_data) != NULL ||
       hypre_ParAMGDataParticipate(amg_data) || hypre_ParAMGDataPartialCycleCoarsestLevel(amg_data) >= 0)
   {
      b200_ctx_sync(ctx);
      fprintf(stderr, "hypre_b200: BoomerAMG option outside the device path (block mode, complex "
              "smoothers, additive cycles, AIR restriction, user-defined relax points, sequential coarse solve)\n");
      return NULL;
   }
   for (k = 0; k < 4; k++) { relax_type[k] = grt[k]; num_sweeps[k] = ngs[k]; }
   if (nl == 1)
   {
      relax_type[0] = hypre_ParAMGDataUserRelaxType(amg_data);
      if (relax_type[0] == -1) { relax_type[0] = 6; }
   }
   if (relax_type[3] == 19) { relax_type[3] = 9; }
   /* relax_type 0 (weighted point Jacobi, hypre_BoomerAMGRelax0WeightedJacobi, par_relax.c:179-335:
    * u_i = (1-w) u_i + w (f_i - sum_{j != i} a_ij u_j) / a_ii) is u + w (f - A u) ./ d with d = the diagonal: the
    * device's Jacobi kernel (relax 7) with the diagonal in the place of the l1 norms.  The setup holds no l1 norms
    * for it; mixing it with the l1 smoothers (one scaling vector per level on the device) is rejected. */
   for (k = 0; k < 4; k++) { if (relax_type[k] == 0) { uses_jac0 = 1; } }
   if (uses_jac0)
   {
      for (k = 0; k < 4; k++)
      {
         if (relax_type[k] == 7 || relax_type[k] == 18)
         {
            b200_ctx_sync(ctx);
            fprintf(stderr, "hypre_b200: relax_type 0 cannot be combined with relax_type 7 / 18 in one cycle on the device path\n");
            return NULL;
         }
         if (relax_type[k] == 0) { relax_type[k] = 7; }
      }
   }

   if (g_nranks > 1 && !g_native)
   {
      /* Row partition of every level.  Level 0: hypre's even split.  Level l+1 inherits the
       * partition of level l: coarse point c belongs to the rank that owns its fine point
       * (C points keep their relative order, so the ranges stay contiguous) -- the layout the
       * reference's parallel coarsening produces (coarse rows stay with their owners). */
      hypre_IntArray **cfm = hypre_ParAMGDataCFMarkerArray(amg_data);
      starts = (int *) calloc((size_t)(nl + 1) * (g_nranks + 1), sizeof(int));
      even_starts((int) hypre_ParCSRMatrixGlobalNumRows(A_array[0]), g_nranks, STARTS(0));
      for (l = 0; l < nl - 1; l++)
      {
         int nf = (int) hypre_ParCSRMatrixGlobalNumRows(A_array[l]);
         int nc = (int) hypre_ParCSRMatrixGlobalNumRows(A_array[l + 1]);
         int ok = 0;
         if (cfm && cfm[l] && hypre_IntArraySize(cfm[l]) == nf)
         {
            const HYPRE_Int *m = hypre_IntArrayData(cfm[l]);
            int i, r = 0, cnt = 0;
            for (i = 0; i < nf; i++)
            {
               while (r < g_nranks && i == STARTS(l)[r]) { STARTS(l + 1)[r++] = cnt; }
               if (m[i] > 0) { cnt++; }
            }
            while (r <= g_nranks) { STARTS(l + 1)[r++] = cnt; }
            ok = (cnt == nc);
         }
         if (!ok) { even_starts(nc, g_nranks, STARTS(l + 1)); }
      }
      /* Small levels are latency-bound: below B200_REPL_NNZ nonzeros (default 2e6; measured at 8 GPUs:
       * 15.2 ms per 256^3 solve against 16.5 ms with 8e6, profiles/r02_bench_c3_8gpu*.json) every rank keeps the whole level and runs it without communication -- the
       * reference's own answer to idle ranks on coarse grids is seq_threshold / hypre_seqAMGCycle
       * (par_cycle.c:362-369).  Results do not depend on this choice for Jacobi / Chebyshev. */
      {
         const char *e = getenv("B200_REPL_NNZ");
         double limit = e ? atof(e) : 2.0e6;
         for (l = 1; l < nl; l++)
         {
            hypre_CSRMatrix *d = hypre_ParCSRMatrixDiag(A_array[l]);
            if ((double) hypre_CSRMatrixNumNonzeros(d) <= limit && hypre_CSRMatrixNumRows(d) <= (1 << 18)) { repl = l; break; }
         }
      }
   }
#ifndef HYPRE_SEQUENTIAL
   if (g_nranks > 1 && g_native)
   {
      /* same rule as the replicated-setup mode, on the GLOBAL size of the level */
      const char *e = getenv("B200_REPL_NNZ");
      double limit = e ? atof(e) : 2.0e6;
      for (l = 1; l < nl && limit > 0.0; l++)
      {
         hypre_CSRMatrix *d = hypre_ParCSRMatrixDiag(A_array[l]), *o = hypre_ParCSRMatrixOffd(A_array[l]);
         HYPRE_Real loc = (HYPRE_Real) hypre_CSRMatrixNumNonzeros(d) + (HYPRE_Real) hypre_CSRMatrixNumNonzeros(o), all = 0.0;
         hypre_MPI_Allreduce(&loc, &all, 1, HYPRE_MPI_REAL, hypre_MPI_SUM, hypre_ParCSRMatrixComm(A_array[l]));
         if (all <= limit && hypre_ParCSRMatrixGlobalNumRows(A_array[l]) <= (1 << 18)) { repl = l; break; }
      }
      rows_of = (int *) calloc((size_t) nl * g_nranks + 1, sizeof(int));
   }
#endif
   if (b200_amg_create(ctx, nl, &amg) != B200_OK) { free(starts); free(rows_of); return NULL; }
   for (l = 0; l < nl; l++)
   {
      b200_parcsr *dA, *dP;
      if (g_native)
      {
#ifndef HYPRE_SEQUENTIAL
         if (l >= repl)
         {
            /* replicated level: gathered on every rank; two-stage GS keeps the owners' blocks (the reference's
             * lower triangles are those of the ranks' diag blocks, par_relax.c:1505-1588) */
            dA = upload_parcsr_gathered(A_array[l], rows_of + (size_t) l * g_nranks);
            if (dA) { b200_parcsr_set_gs_blocks(dA, g_nranks, rows_of + (size_t) l * g_nranks); }
            dP = (l < nl - 1) ? upload_parcsr_gathered(P_array[l], NULL) : NULL;
         }
         else if (l + 1 == repl)
         {
            dA = upload_parcsr(A_array[l]);
            dP = upload_parcsr_local_slab(P_array[l]);
         }
         else
#endif
         {
            /* hypre's own distributed objects: what this rank owns, with the CommPkg hypre negotiated */
            dA = upload_parcsr(A_array[l]);
            dP = (l < nl - 1) ? upload_parcsr(P_array[l]) : NULL;
         }
      }
      else if (g_nranks > 1 && l >= repl)
      {
         /* replicated level: the whole single-rank matrices */
         dA = upload_parcsr(A_array[l]);
         dP = (l < nl - 1) ? upload_parcsr(P_array[l]) : NULL;
      }
      else if (g_nranks > 1)
      {
         dA = upload_parcsr_rows(A_array[l], STARTS(l), STARTS(l));
         if (l == nl - 1) { dP = NULL; }
         else if (l + 1 == repl) { dP = upload_parcsr_row_slab(P_array[l], STARTS(l)[g_rank], STARTS(l)[g_rank + 1]); }
         else { dP = upload_parcsr_rows(P_array[l], STARTS(l), STARTS(l + 1)); }
      }
      else
      {
         dA = upload_parcsr(A_array[l]);
         dP = (l < nl - 1) ? upload_parcsr(P_array[l]) : NULL;
      }
      if (!dA || (l < nl - 1 && !dP) || b200_amg_set_level(amg, l, dA, dP) != B200_OK)
      {
         if (dA) { b200_parcsr_destroy(dA); }
         if (dP) { b200_parcsr_destroy(dP); }
         b200_amg_destroy(amg);
         free(starts); free(rows_of);
         return NULL;
      }
   }
   if (b200_amg_set_cycle(amg, relax_type, num_sweeps, hypre_ParAMGDataCycleType(amg_data),
                          hypre_ParAMGDataFCycle(amg_data)) != B200_OK) { b200_amg_destroy(amg); free(starts); return NULL; }
   if (repl < nl && b200_amg_set_replicated_from(amg, repl) != B200_OK) { b200_amg_destroy(amg); free(starts); return NULL; }
   /* CF-ordered relaxation (HYPRE_BoomerAMGSetRelaxOrder(1), ij -CF 1): the CF markers of the setup, this rank's rows */
   if (hypre_ParAMGDataRelaxOrder(amg_data) == 1 && nl > 1)
   {
      hypre_IntArray **cfm = hypre_ParAMGDataCFMarkerArray(amg_data);
      int st = b200_amg_set_relax_order(amg, 1);
      for (l = 0; l < nl - 1 && st == B200_OK; l++)
      {
         const int off = (starts && l < repl) ? STARTS(l)[g_rank] : 0;
         const HYPRE_Int n_here = hypre_CSRMatrixNumRows(hypre_ParCSRMatrixDiag(A_array[l]));
         const int *cf_here = (cfm && cfm[l]) ? hypre_IntArrayData(cfm[l]) : NULL;
         if (cf_here == NULL && n_here > 0) { st = B200_ERR_ARG; break; }       /* (a rank without rows holds no marker) */
         if (g_native && rows_of != NULL && l >= repl)
         {
#ifndef HYPRE_SEQUENTIAL
            /* gathered level: every rank's markers, in rank order */
            const int *ro = rows_of + (size_t) l * g_nranks;
            int r, *dsp = (int *) calloc((size_t) g_nranks + 1, sizeof(int)), *all;
            for (r = 0; r < g_nranks; r++) { dsp[r + 1] = dsp[r] + ro[r]; }
            all = (int *) calloc((size_t) (dsp[g_nranks] > 0 ? dsp[g_nranks] : 1), sizeof(int));
            int dummy = 0;
            hypre_MPI_Allgatherv((void *) (cf_here ? cf_here : &dummy), ro[g_rank], HYPRE_MPI_INT, all, (HYPRE_Int *) ro, dsp, HYPRE_MPI_INT,
                                 hypre_ParCSRMatrixComm(A_array[l]));
            st = b200_amg_set_level_cf_marker(amg, l, all);
            free(dsp); free(all);
#endif
         }
         else { st = b200_amg_set_level_cf_marker(amg, l, cf_here ? cf_here + off : NULL); }
      }
      if (st != B200_OK)
      {
         fprintf(stderr, "hypre_b200: relax_order 1 without CF markers on every level\n");
         b200_amg_destroy(amg); free(starts); free(rows_of); return NULL;
      }
   }
   for (k = 0; k < 4; k++) { if (relax_type[k] == 16) { uses_cheby = 1; } }
   for (l = 0; l < nl; l++)
   {
      const int off = (starts && l < repl) ? STARTS(l)[g_rank] : 0;   /* this rank's slice of the level vectors */
      const double *l1d = (l1 && l1[l]) ? hypre_VectorData(l1[l]) + off : NULL;
      double *diag0 = NULL, *l1_all = NULL, *ds_all = NULL;
      int status;
      const int gathered = g_native && rows_of != NULL && l >= repl;
      (void) gathered;
      if (uses_jac0)
      {
         /* the diagonal of this level (first entry of each row of the diag block) in the place of the l1 norms */
         hypre_CSRMatrix *d = hypre_ParCSRMatrixDiag(A_array[l]);
         HYPRE_Int nr = hypre_CSRMatrixNumRows(d), r;
         diag0 = (double *) calloc((size_t) (nr > 0 ? nr : 1), sizeof(double));
         for (r = 0; r < nr; r++)
         {
            diag0[r] = hypre_CSRMatrixI(d)[r + 1] > hypre_CSRMatrixI(d)[r] ? hypre_CSRMatrixData(d)[hypre_CSRMatrixI(d)[r]] : 0.0;
         }
         l1d = diag0 + off;
      }
#ifndef HYPRE_SEQUENTIAL
      if (gathered)
      {
         /* collective: the decision must not depend on this rank's data (a rank without rows on this level holds no
          * l1 norms at all) */
         HYPRE_Int have = (l1d != NULL || uses_jac0) ? 1 : 0, any = 0;
         hypre_MPI_Allreduce(&have, &any, 1, HYPRE_MPI_INT, hypre_MPI_MAX, hypre_ParCSRMatrixComm(A_array[l]));
         if (any)
         {
            l1_all = gather_level_vector(hypre_ParCSRMatrixComm(A_array[l]), l1d, rows_of + (size_t) l * g_nranks);
            l1d = l1_all;
         }
      }
#endif
      status = b200_amg_set_level_relax(amg, l, hypre_ParAMGDataRelaxWeight(amg_data)[l], l1d);
      free(diag0); free(l1_all);
      if (status != B200_OK)
      {
         b200_amg_destroy(amg); free(starts); return NULL;
      }
      if (uses_cheby && coefs && coefs[l])
      {
         /* unscaled Chebyshev (cheby_scale 0) has no scaling vector: the setup leaves ds[l] without data */
         const double *dsd = (hypre_ParAMGDataChebyScale(amg_data) && ds && ds[l] && hypre_VectorData(ds[l]))
                             ? hypre_VectorData(ds[l]) + off : NULL;
#ifndef HYPRE_SEQUENTIAL
         if (gathered && hypre_ParAMGDataChebyScale(amg_data))       /* (same decision on every rank) */
         {
            ds_all = gather_level_vector(hypre_ParCSRMatrixComm(A_array[l]), dsd, rows_of + (size_t) l * g_nranks);
            dsd = ds_all;
         }
#endif
         status = b200_amg_set_level_cheby(amg, l, dsd, coefs[l], hypre_ParAMGDataChebyOrder(amg_data),
                                           hypre_ParAMGDataChebyScale(amg_data));
         free(ds_all);
         if (status != B200_OK)
         {
            b200_amg_destroy(amg); free(starts); return NULL;
         }
      }
   }
   if (relax_type[3] == 9 && nl > 1)
   {
      hypre_ParCSRMatrix *Ac = A_array[nl - 1];
      HYPRE_Int N = (HYPRE_Int) hypre_ParCSRMatrixGlobalNumRows(Ac);
      HYPRE_Int np;
      int status;
      if (!hypre_ParAMGDataGSSetup(amg_data)) { hypre_GaussElimSetup(amg_data, nl - 1, grt[3]); }
      hypre_MPI_Comm_size(hypre_ParCSRMatrixComm(Ac), &np);
#ifndef HYPRE_SEQUENTIAL
      if (g_native && repl < nl)
      {
         /* the coarsest level is replicated: every rank holds the whole dense matrix (gathered from the owners' rows)
          * and solves locally -- no gather of the right-hand side per cycle */
         MPI_Comm comm = hypre_ParCSRMatrixComm(Ac);
         int *li, *lj, r, k, mine = (int) hypre_CSRMatrixNumRows(hypre_ParCSRMatrixDiag(Ac));
         double *la, *dense_loc = (double *) calloc((size_t) (mine > 0 ? mine : 1) * (size_t) N, sizeof(double));
         double *dense = (double *) calloc((size_t) N * (size_t) N, sizeof(double));
         int *cnt = (int *) calloc((size_t) g_nranks, sizeof(int)), *dsp = (int *) calloc((size_t) g_nranks + 1, sizeof(int));
         int *rows = (int *) calloc((size_t) g_nranks, sizeof(int));
         int sz = mine * (int) N;
         local_rows_global_cols(Ac, &li, &lj, &la);
         for (r = 0; r < mine; r++) { for (k = li[r]; k < li[r + 1]; k++) { dense_loc[(size_t) r * N + lj[k]] = la[k]; } }
         hypre_MPI_Allgather(&sz, 1, HYPRE_MPI_INT, cnt, 1, HYPRE_MPI_INT, comm);
         for (r = 0; r < g_nranks; r++) { dsp[r + 1] = dsp[r] + cnt[r]; rows[r] = (int) N; }
         hypre_MPI_Allgatherv(dense_loc, sz, HYPRE_MPI_REAL, dense, cnt, dsp, HYPRE_MPI_REAL, comm);
         status = b200_amg_set_coarse_dense(amg, N, dense, 0, rows);
         free(li); free(lj); free(la); free(dense_loc); free(dense); free(cnt); free(dsp); free(rows);
      }
      else
#endif
      if (g_native)
      {
         /* rows of the coarsest level per WORLD rank (hypre's comm_info is indexed by the ranks of the
          * sub-communicator of owners, par_gauss_elim.c:60-100); ranks without coarse rows hold no A_mat */
         HYPRE_Int mine = hypre_CSRMatrixNumRows(hypre_ParCSRMatrixDiag(Ac));
         HYPRE_Int *rows = (HYPRE_Int *) calloc((size_t) g_nranks, sizeof(HYPRE_Int));
         hypre_MPI_Allgather(&mine, 1, HYPRE_MPI_INT, rows, 1, HYPRE_MPI_INT, hypre_ParCSRMatrixComm(Ac));
         status = b200_amg_set_coarse_dense(amg, N, mine > 0 ? hypre_ParAMGDataAMat(amg_data) : NULL,
                                            (int) hypre_ParCSRMatrixFirstRowIndex(Ac), rows);
         free(rows);
      }
      else if (starts)
      {
         int *rows = (int *) calloc((size_t) g_nranks, sizeof(int)), r;
         for (r = 0; r < g_nranks; r++) { rows[r] = repl < nl ? (int) N : STARTS(nl - 1)[r + 1] - STARTS(nl - 1)[r]; }
         status = b200_amg_set_coarse_dense(amg, N, hypre_ParAMGDataAMat(amg_data),
                                            repl < nl ? 0 : STARTS(nl - 1)[g_rank], rows);
         free(rows);
      }
      else
      {
         status = b200_amg_set_coarse_dense(amg, N, hypre_ParAMGDataAMat(amg_data),
                                            (int) hypre_ParCSRMatrixFirstRowIndex(Ac),
                                            np > 1 ? hypre_ParAMGDataCommInfo(amg_data) : NULL);
      }
      if (status != B200_OK) { b200_amg_destroy(amg); free(starts); return NULL; }
   }
   if (starts_out) { *starts_out = starts; starts = NULL; }
   if (repl_out) { *repl_out = repl; }
   free(starts); free(rows_of);
#undef STARTS
   if (b200_amg_finalize(amg) != B200_OK) { b200_amg_destroy(amg); return NULL; }
   if (verbose())
   {
      double bytes = 0.0;
      b200_amg_cycle_bytes(amg, &bytes);
      fprintf(stderr, "hypre_b200: rank %d/%d uploaded %d-level hierarchy (replicated from level %d), %.3f MB algorithmic bytes per cycle\n",
              g_rank, g_nranks, (int) nl, repl < nl ? repl : -1, bytes * 1e-6);
   }
   return amg;
}

/* --------------------------------------------------------- interposition */
typedef HYPRE_Int (*pcg_solve_fn)(void *, void *, void *, void *);
typedef HYPRE_Int (*amg_setup_fn)(void *, hypre_ParCSRMatrix *, hypre_ParVector *, hypre_ParVector *);
typedef HYPRE_Int (*amg_destroy_fn)(void *);
typedef HYPRE_Int (*mat_destroy_fn)(hypre_ParCSRMatrix *);

static void *next_sym(const char *name)
{
   void *p = dlsym(RTLD_NEXT, name);
   if (!p) { fprintf(stderr, "hypre_b200: reference symbol %s not found\n", name); abort(); }
   return p;
}

/* A new setup invalidates the device mirror of that solver's hierarchy. */
HYPRE_Int hypre_BoomerAMGSetup(void *amg_vdata, hypre_ParCSRMatrix *A, hypre_ParVector *f, hypre_ParVector *u)
{
   static amg_setup_fn next = NULL;
   if (!next) { next = (amg_setup_fn) next_sym("hypre_BoomerAMGSetup"); }
   mirror_drop(amg_vdata);
   return next(amg_vdata, A, f, u);
}

HYPRE_Int hypre_BoomerAMGDestroy(void *amg_vdata)
{
   static amg_destroy_fn next = NULL;
   if (!next) { next = (amg_destroy_fn) next_sym("hypre_BoomerAMGDestroy"); }
   mirror_drop(amg_vdata);
   return next(amg_vdata);
}

/* a vector that goes away takes its device binding with it (its address may be handed out again) */
HYPRE_Int hypre_ParVectorDestroy(hypre_ParVector *v)
{
   typedef HYPRE_Int (*vec_destroy_fn)(hypre_ParVector *);
   static vec_destroy_fn next = NULL;
   if (!next) { next = (vec_destroy_fn) next_sym("hypre_ParVectorDestroy"); }
   if (v && g_devvecs) { hypre_b200_vector_bind_device(v, NULL); }
   return next(v);
}

HYPRE_Int hypre_ParCSRMatrixDestroy(hypre_ParCSRMatrix *A)
{
   static mat_destroy_fn next = NULL;
   if (!next) { next = (mat_destroy_fn) next_sym("hypre_ParCSRMatrixDestroy"); }
   if (A) { mirror_drop(A); }
   return next(A);
}

/* the coefficients of a host matrix changed: drop every device copy made from it (the next solve uploads again;
 * a hierarchy is re-uploaded from the host hierarchy as it is -- new A on level 0, lagged coarse levels -- which
 * is what the reference's CPU solve would use) */
void hypre_b200_matrix_changed(void *parcsr)
{
   int again = 1;
   while (again)
   {
      mirror_t *m;
      again = 0;
      for (m = g_mirrors; m; m = m->next)
      {
         if (m->src == (hypre_ParCSRMatrix *) parcsr) { mirror_drop(m->key); again = 1; break; }
      }
   }
}

static mirror_t *mirror_current(void *key)
{
   mirror_t *m = mirror_find(key);
   if (m && m->src && m->stamp != matrix_stamp(m->src))
   {
      if (verbose()) { fprintf(stderr, "hypre_b200: coefficients of matrix %p changed, device copy dropped\n", (void *) m->src); }
      mirror_drop(key);
      m = NULL;
   }
   return m;
}

/* HYPRE_IJMatrixAssemble (src/IJ_mv/HYPRE_IJMatrix.c) closes a round of SetValues / AddToValues: whatever was
 * uploaded from that matrix is stale */
HYPRE_Int HYPRE_IJMatrixAssemble(HYPRE_IJMatrix matrix)
{
   typedef HYPRE_Int (*ij_assemble_fn)(HYPRE_IJMatrix);
   static ij_assemble_fn next = NULL;
   HYPRE_Int ierr;
   if (!next) { next = (ij_assemble_fn) next_sym("HYPRE_IJMatrixAssemble"); }
   ierr = next(matrix);
   if (matrix && hypre_IJMatrixObjectType((hypre_IJMatrix *) matrix) == HYPRE_PARCSR && hypre_IJMatrixObject((hypre_IJMatrix *) matrix))
   {
      hypre_b200_matrix_changed(hypre_IJMatrixObject((hypre_IJMatrix *) matrix));
   }
   return ierr;
}

static mirror_t *mirror_for_amg(hypre_ParAMGData *amg_data)
{
   mirror_t *m = mirror_current(amg_data);
   if (m) { return m; }
   {
      int *starts = NULL, repl = 1 << 30;
      b200_amg *amg = upload_amg(amg_data, &starts, &repl);
      if (!amg) { free(starts); return NULL; }
      m = (mirror_t *) calloc(1, sizeof(mirror_t));
      m->key = amg_data; m->amg = amg;
      m->src = hypre_ParAMGDataAArray(amg_data)[0]; m->stamp = matrix_stamp(m->src);
      m->starts = starts; m->repl_level = repl; m->num_levels = (int) hypre_ParAMGDataNumLevels(amg_data);
      b200_amg_level_matrix(amg, 0, &m->A);
      if (g_nranks > 1 && !g_native)
      {
         int *st = (int *) calloc((size_t) g_nranks + 1, sizeof(int));
         even_starts((int) hypre_ParCSRMatrixGlobalNumRows(hypre_ParAMGDataAArray(amg_data)[0]), g_nranks, st);
         m->first_row = m->first_col = st[g_rank];
         free(st);
      }
      m->next = g_mirrors; g_mirrors = m;
   }
   return m;
}

static mirror_t *mirror_for_matrix(hypre_ParCSRMatrix *A)
{
   mirror_t *m = mirror_current(A);
   if (m) { return m; }
   {
      b200_parcsr *dA = NULL;
      int first_row = 0, first_col = 0;
      if (glue_ctx() && g_nranks > 1 && !g_native)
      {
         int *rs = (int *) calloc((size_t) g_nranks + 1, sizeof(int));
         int *cs = (int *) calloc((size_t) g_nranks + 1, sizeof(int));
         even_starts((int) hypre_ParCSRMatrixGlobalNumRows(A), g_nranks, rs);
         even_starts((int) hypre_ParCSRMatrixGlobalNumCols(A), g_nranks, cs);
         dA = upload_parcsr_rows(A, rs, cs);
         first_row = rs[g_rank]; first_col = cs[g_rank];
         free(rs); free(cs);
      }
      else if (glue_ctx()) { dA = upload_parcsr(A); }
      if (!dA) { return NULL; }
      m = (mirror_t *) calloc(1, sizeof(mirror_t));
      m->key = A; m->A = dA; m->first_row = first_row; m->first_col = first_col;
      m->src = A; m->stamp = matrix_stamp(A);
      m->next = g_mirrors; g_mirrors = m;
   }
   return m;
}

/* hypre_PCGSolve (src/krylov/pcg.c:317-1008) for the ParCSR operator table. */
HYPRE_Int hypre_PCGSolve(void *pcg_vdata, void *A, void *b, void *x)
{
   static pcg_solve_fn next = NULL;
   hypre_PCGData      *pcg = (hypre_PCGData *) pcg_vdata;
   hypre_PCGFunctions *fn  = pcg->functions;
   hypre_ParCSRMatrix *pA  = (hypre_ParCSRMatrix *) A;
   hypre_ParVector    *pb  = (hypre_ParVector *) b, *px = (hypre_ParVector *) x;
   HYPRE_Int (*precond)(void *, void *, void *, void *) = fn->precond;
   mirror_t *m = NULL;
   b200_amg *amg = NULL;
   b200_pcg_params prm;
   b200_pcg_result res;
   HYPRE_Int my_id, num_procs, i;
   int status;

   if (!next) { next = (pcg_solve_fn) next_sym("hypre_PCGSolve"); }
   if (!g_enabled) { return next(pcg_vdata, A, b, x); }
   /* other interfaces (struct, sstruct, ...) were never on this path */
   if (fn->Matvec != hypre_ParKrylovMatvec) { return next(pcg_vdata, A, b, x); }

   memset(&prm, 0, sizeof(prm));
   if (pcg->rel_change || pcg->recompute_residual || pcg->recompute_residual_p || pcg->stop_crit ||
       pcg->flex || pcg->rtol != 0.0 || pcg->cf_tol > 0.0 || pcg->atolf > 0.0 ||
       hypre_ParVectorNumVectors(pb) > 1)
   {
      GLUE_FAIL("PCG option outside the device path (rel_change / recompute_residual / stop_crit / flex / rtol / cf_tol / atolf / multivector)");
      return hypre_error_flag;
   }
   if (precond == (HYPRE_Int (*)(void *, void *, void *, void *)) HYPRE_BoomerAMGSolve ||
       precond == (HYPRE_Int (*)(void *, void *, void *, void *)) hypre_BoomerAMGSolve)
   {
      hypre_ParAMGData *amg_data = (hypre_ParAMGData *) pcg->precond_data;
      if (hypre_ParAMGDataTol(amg_data) != 0.0 || hypre_ParAMGDataMaxIter(amg_data) < 1 || hypre_ParAMGDataMinIter(amg_data) > hypre_ParAMGDataMaxIter(amg_data))
      {
         GLUE_FAIL("BoomerAMG as preconditioner must use tol 0 (a fixed number of cycles, max_iter >= 1)");
         return hypre_error_flag;
      }
      if (hypre_ParAMGDataAArray(amg_data) == NULL || hypre_ParAMGDataAArray(amg_data)[0] != pA)
      {
         GLUE_FAIL("BoomerAMG was not set up for this matrix");
         return hypre_error_flag;
      }
      m = mirror_for_amg(amg_data);
      if (!m) { GLUE_FAIL("hierarchy upload failed"); return hypre_error_flag; }
      amg = m->amg;
      b200_amg_set_precond_cycles(amg, (int) hypre_ParAMGDataMaxIter(amg_data));     /* ij -precon_cycles */
   }
   else if (precond == (HYPRE_Int (*)(void *, void *, void *, void *)) HYPRE_ParCSRDiagScale)
   {
      prm.diag_scale = 1;
      m = mirror_for_matrix(pA);
   }
   else if (precond == (HYPRE_Int (*)(void *, void *, void *, void *)) hypre_ParKrylovIdentity)
   {
      m = mirror_for_matrix(pA);
   }
   else
   {
      GLUE_FAIL("preconditioner has no device implementation (BoomerAMG, diagonal scaling, identity are supported)");
      return hypre_error_flag;
   }
   if (!m) { GLUE_FAIL("matrix upload failed"); return hypre_error_flag; }

   prm.tol = pcg->tol; prm.a_tol = pcg->a_tol; prm.max_iter = pcg->max_iter;
   prm.two_norm = pcg->two_norm; prm.skip_break = pcg->skip_break;
   pcg->converged = 0;
   (*(fn->CommInfo))(A, &my_id, &num_procs);

   /* distributed mode: this rank's slice of the replicated host vectors (offset 0 otherwise) */
   {
      double *db, *dx;
      double *hn = (pcg->logging > 0 || pcg->print_level > 0) ? pcg->norms : NULL;
      double *hr = (pcg->logging > 0 || pcg->print_level > 0) ? pcg->rel_norms : NULL;
      if (krylov_operands(pb, px, &db, &dx) != 0) { return hypre_error_flag; }
      /* device-bound vectors: solve in place on the application's arrays (no copy of b or x) */
      status = db ? b200_pcg_solve(glue_ctx(), m->A, amg, &prm, db, dx, &res, hn, hr)
                  : b200_pcg_solve_host(glue_ctx(), m->A, amg, &prm,
                                        hypre_VectorData(hypre_ParVectorLocalVector(pb)) + m->first_row,
                                        hypre_VectorData(hypre_ParVectorLocalVector(px)) + m->first_row, &res, hn, hr);
   }
   if (status != B200_OK) { GLUE_FAIL("device PCG failed"); return hypre_error_flag; }
   if (verbose() && g_rank == 0)
   {
      fprintf(stderr, "hypre_b200: PCG solve on the device: %d iterations, relative residual %.6e\n",
              res.num_iterations, res.rel_residual_norm);
   }

   /* the reference's printed table (pcg.c:405-409,536-560,772-792,983-986) */
   if (pcg->print_level > 1 && my_id == 0 && g_rank == 0)
   {
      hypre_printf(pcg->two_norm ? "<b,b>: %e\n" : "<C*b,b>: %e\n", res.bi_prod);
      if (res.error_flag != 1 && res.bi_prod > 0.0)
      {
         hypre_printf("\n\n");
         if (pcg->two_norm)
         {
            hypre_printf("Iters       ||r||_2     conv.rate  ||r||_2/||b||_2\n");
            hypre_printf("-----    ------------   ---------  ------------ \n");
         }
         else
         {
            hypre_printf("Iters       ||r||_C     conv.rate  ||r||_C/||b||_C\n");
            hypre_printf("-----    ------------    ---------  ------------ \n");
         }
         for (i = 1; i <= res.num_iterations; i++)
         {
            hypre_printf("% 5d    %e    %f    %e\n", i, pcg->norms[i], pcg->norms[i] / pcg->norms[i - 1],
                         pcg->rel_norms[i]);
         }
         hypre_printf("\n\n");
      }
   }
   if (res.error_flag == 1)
   {
      if (pcg->print_level > 0 || pcg->logging > 0)
      {
         hypre_printf("\n\nERROR detected by Hypre ...  BEGIN\n");
         hypre_printf("ERROR -- hypre_PCGSolve: INFs and/or NaNs detected in input.\n");
         hypre_printf("Returning error flag += 101.  Program not terminated.\n");
         hypre_printf("ERROR detected by Hypre ...  END\n\n\n");
      }
      hypre_error(HYPRE_ERROR_GENERIC);
      return hypre_error_flag;
   }
   if (res.error_flag & 256) { hypre_error_w_msg(HYPRE_ERROR_CONV, "PCG (B200 path): breakdown or max iterations reached"); }
   pcg->converged = res.converged;
   pcg->num_iterations = res.num_iterations;
   pcg->rel_residual_norm = res.rel_residual_norm;
   hypre_ParVectorAllZeros(px) = 0;
   return hypre_error_flag;
}

/* ------------------------------------------------------------ GMRES / BiCGSTAB
 * Preconditioner of a ParCSR Krylov solver -> device objects: BoomerAMG (one cycle), diagonal scaling or
 * identity, as for PCG above.  Returns NULL after GLUE_FAIL when there is no device implementation. */
static mirror_t *krylov_precond_mirror(HYPRE_Int (*precond)(void *, void *, void *, void *), void *precond_data,
                                       hypre_ParCSRMatrix *pA, b200_amg **amg, int *diag_scale)
{
   mirror_t *m = NULL;
   *amg = NULL; *diag_scale = 0;
   if (precond == (HYPRE_Int (*)(void *, void *, void *, void *)) HYPRE_BoomerAMGSolve ||
       precond == (HYPRE_Int (*)(void *, void *, void *, void *)) hypre_BoomerAMGSolve)
   {
      hypre_ParAMGData *amg_data = (hypre_ParAMGData *) precond_data;
      if (hypre_ParAMGDataTol(amg_data) != 0.0 || hypre_ParAMGDataMaxIter(amg_data) < 1 || hypre_ParAMGDataMinIter(amg_data) > hypre_ParAMGDataMaxIter(amg_data))
      {
         GLUE_FAIL("BoomerAMG as preconditioner must use tol 0 (a fixed number of cycles, max_iter >= 1)");
         return NULL;
      }
      if (hypre_ParAMGDataAArray(amg_data) == NULL || hypre_ParAMGDataAArray(amg_data)[0] != pA)
      {
         GLUE_FAIL("BoomerAMG was not set up for this matrix");
         return NULL;
      }
      m = mirror_for_amg(amg_data);
      if (!m) { GLUE_FAIL("hierarchy upload failed"); return NULL; }
      *amg = m->amg;
      b200_amg_set_precond_cycles(*amg, (int) hypre_ParAMGDataMaxIter(amg_data));    /* ij -precon_cycles */
   }
   else if (precond == (HYPRE_Int (*)(void *, void *, void *, void *)) HYPRE_ParCSRDiagScale)
   {
      *diag_scale = 1;
      m = mirror_for_matrix(pA);
      if (!m) { GLUE_FAIL("matrix upload failed"); }
   }
   else if (precond == (HYPRE_Int (*)(void *, void *, void *, void *)) hypre_ParKrylovIdentity)
   {
      m = mirror_for_matrix(pA);
      if (!m) { GLUE_FAIL("matrix upload failed"); }
   }
   else
   {
      GLUE_FAIL("preconditioner has no device implementation (BoomerAMG, diagonal scaling, identity are supported)");
   }
   return m;
}

static void krylov_nan_message(const char *who)
{
   hypre_printf("\n\nERROR detected by Hypre ... BEGIN\n");
   hypre_printf("ERROR -- %s: INFs and/or NaNs detected in input.\n", who);
   hypre_printf("Returning error flag += 101.  Program not terminated.\n");
   hypre_printf("ERROR detected by Hypre ... END\n\n\n");
}

/* residual table of hypre_GMRESSolve / hypre_BiCGSTABSolve (gmres.c:417-460, 612-630; bicgstab.c:354-362,
 * 406-421, 491-499) from the history the device solve returned */
static void krylov_print_table(const double *norms, int iters, double b_norm)
{
   int i;
   hypre_printf("L2 norm of b: %e\n", b_norm);
   if (b_norm == 0.0) { hypre_printf("Rel_resid_norm actually contains the residual norm\n"); }
   hypre_printf("Initial L2 norm of residual: %e\n", norms[0]);
   hypre_printf("=============================================\n\n");
   if (b_norm > 0.0)
   {
      hypre_printf("Iters     resid.norm     conv.rate  rel.res.norm\n");
      hypre_printf("-----    ------------    ---------- ------------\n");
   }
   else
   {
      hypre_printf("Iters     resid.norm     conv.rate\n");
      hypre_printf("-----    ------------    ----------\n");
   }
   for (i = 1; i <= iters; i++)
   {
      if (b_norm > 0.0) { hypre_printf("% 5d    %e    %f   %e\n", i, norms[i], norms[i] / norms[i - 1], norms[i] / b_norm); }
      else { hypre_printf("% 5d    %e    %f\n", i, norms[i], norms[i] / norms[i - 1]); }
   }
}

/* Everything hypre_GMRESSolve (kind 0, print_level > 1) / hypre_COGMRESSolve (kind 2: as GMRES without the closing
 * blank lines) / hypre_BiCGSTABSolve (kind 1, print_level > 0) print
 * for a solve with this history: table, "Final L2 norm of residual" after the exits that evaluated b - A x
 * (gmres.c:512-520, 783-851, 895-898; bicgstab.c:436-441, 508-512).  Host printing only; exported so that the
 * CPU tests can compare the text with the reference driver's. */
void hypre_b200_krylov_report(int kind, const double *norms, int iters, double b_norm, double rel_norm, int real_residual_exit)
{
   krylov_print_table(norms, iters, b_norm);
   if (real_residual_exit)
   {
      hypre_printf("\n\n");
      if (kind == 1 && iters == 0) { hypre_printf("Tolerance and min_iter requirements satisfied by initial data.\n"); }
      hypre_printf("Final L2 norm of residual: %e\n\n", b_norm > 0.0 ? rel_norm * b_norm : rel_norm);
   }
   if (kind == 0) { hypre_printf("\n\n"); }
}

/* hypre_GMRESSolve (src/krylov/gmres.c:274-930) for the ParCSR operator table. */
HYPRE_Int hypre_GMRESSolve(void *gmres_vdata, void *A, void *b, void *x)
{
   static pcg_solve_fn next = NULL;
   hypre_GMRESData      *g  = (hypre_GMRESData *) gmres_vdata;
   hypre_GMRESFunctions *fn = g->functions;
   hypre_ParCSRMatrix   *pA = (hypre_ParCSRMatrix *) A;
   hypre_ParVector      *pb = (hypre_ParVector *) b, *px = (hypre_ParVector *) x;
   mirror_t *m;
   b200_amg *amg = NULL;
   b200_gmres_params prm;
   b200_krylov_result res;
   HYPRE_Int my_id, num_procs;
   double *norms, *own_norms = NULL;
   int status, diag_scale = 0;

   if (!next) { next = (pcg_solve_fn) next_sym("hypre_GMRESSolve"); }
   if (!g_enabled) { return next(gmres_vdata, A, b, x); }
   if (fn->Matvec != hypre_ParKrylovMatvec) { return next(gmres_vdata, A, b, x); }

   /* hybrid == -1 (set by HYPRE_ParCSRHybrid on its inner solver, parcsr_ls/amg_hybrid.c:1968, 2029; krylov/gmres.c:908) only
    * silences HYPRE_ERROR_CONV at the end */
   if (g->rel_change || g->stop_crit || (g->hybrid != 0 && g->hybrid != -1) || hypre_ParVectorNumVectors(pb) > 1)
   {
      GLUE_FAIL("GMRES option outside the device path (rel_change / stop_crit / multivector)");
      return hypre_error_flag;
   }
   m = krylov_precond_mirror((HYPRE_Int (*)(void *, void *, void *, void *)) fn->precond, g->precond_data, pA, &amg, &diag_scale);
   if (!m) { return hypre_error_flag; }

   memset(&prm, 0, sizeof(prm));
   prm.tol = g->tol; prm.a_tol = g->a_tol; prm.max_iter = g->max_iter; prm.min_iter = g->min_iter;
   prm.k_dim = g->k_dim; prm.skip_real_r_check = g->skip_real_r_check; prm.diag_scale = diag_scale;
   prm.cf_tol = g->cf_tol;
   g->converged = 0;
   (*(fn->CommInfo))(A, &my_id, &num_procs);
   /* norms exists when logging or print_level is set (hypre_GMRESSetup, gmres.c:246-256).  The reference
    * fills norms[0] then and norms[iter] only when print_level > 0 (gmres.c:413, 610); here the whole history
    * is stored in both cases */
   norms = ((g->logging > 0 || g->print_level > 0) && g->norms) ? g->norms : NULL;
   if (!norms && g->print_level > 1) { norms = own_norms = (double *) calloc((size_t) g->max_iter + 1, sizeof(double)); }

   {
      double *db, *dx;
      if (krylov_operands(pb, px, &db, &dx) != 0) { free(own_norms); return hypre_error_flag; }
      status = db ? b200_gmres_solve(glue_ctx(), m->A, amg, &prm, db, dx, &res, norms)
                  : b200_gmres_solve_host(glue_ctx(), m->A, amg, &prm,
                                          hypre_VectorData(hypre_ParVectorLocalVector(pb)) + m->first_row,
                                          hypre_VectorData(hypre_ParVectorLocalVector(px)) + m->first_row, &res, norms);
   }
   if (status != B200_OK) { free(own_norms); GLUE_FAIL("device GMRES failed"); return hypre_error_flag; }
   if (verbose() && g_rank == 0)
   {
      fprintf(stderr, "hypre_b200: GMRES(%d) solve on the device: %d iterations, relative residual %.6e\n",
              (int) g->k_dim, res.num_iterations, res.rel_residual_norm);
   }
   if (res.error_flag & 1)
   {
      if (g->logging > 0 || g->print_level > 0) { krylov_nan_message("hypre_GMRESSolve"); }
      free(own_norms);
      hypre_error(HYPRE_ERROR_GENERIC);
      return hypre_error_flag;
   }
   if (g->print_level > 1 && my_id == 0 && g_rank == 0)
   {
      hypre_b200_krylov_report(0, norms, res.num_iterations, res.b_norm, res.rel_residual_norm, res.real_residual_exit);
   }
   free(own_norms);
   g->converged = res.converged;
   g->num_iterations = res.num_iterations;
   g->rel_residual_norm = res.rel_residual_norm;
   hypre_ParVectorAllZeros(px) = 0;
   if ((res.error_flag & 256) && g->hybrid != -1) { hypre_error(HYPRE_ERROR_CONV); }
   return hypre_error_flag;
}

/* hypre_COGMRESSolve (src/krylov/cogmres.c:277-…) for the ParCSR operator table: GMRES with classical
 * Gram-Schmidt through MassInnerProd / MassAxpy (cgs == 1; `unroll` only regroups the host loops). */
HYPRE_Int hypre_COGMRESSolve(void *cogmres_vdata, void *A, void *b, void *x)
{
   static pcg_solve_fn next = NULL;
   hypre_COGMRESData      *g  = (hypre_COGMRESData *) cogmres_vdata;
   hypre_COGMRESFunctions *fn = g->functions;
   hypre_ParCSRMatrix     *pA = (hypre_ParCSRMatrix *) A;
   hypre_ParVector        *pb = (hypre_ParVector *) b, *px = (hypre_ParVector *) x;
   mirror_t *m;
   b200_amg *amg = NULL;
   b200_gmres_params prm;
   b200_krylov_result res;
   HYPRE_Int my_id, num_procs;
   double *norms, *own_norms = NULL;
   int status, diag_scale = 0;

   if (!next) { next = (pcg_solve_fn) next_sym("hypre_COGMRESSolve"); }
   if (!g_enabled) { return next(cogmres_vdata, A, b, x); }
   if (fn->Matvec != hypre_ParKrylovMatvec) { return next(cogmres_vdata, A, b, x); }

   if (g->rel_change || g->cgs > 1 || hypre_ParVectorNumVectors(pb) > 1)
   {
      GLUE_FAIL("COGMRES option outside the device path (rel_change / cgs 2 / multivector)");
      return hypre_error_flag;
   }
   m = krylov_precond_mirror((HYPRE_Int (*)(void *, void *, void *, void *)) fn->precond, g->precond_data, pA, &amg, &diag_scale);
   if (!m) { return hypre_error_flag; }

   memset(&prm, 0, sizeof(prm));
   prm.tol = g->tol; prm.a_tol = g->a_tol; prm.max_iter = g->max_iter; prm.min_iter = g->min_iter;
   prm.k_dim = g->k_dim; prm.skip_real_r_check = g->skip_real_r_check; prm.diag_scale = diag_scale;
   prm.ortho = 1; prm.cf_tol = g->cf_tol;
   g->converged = 0;
   (*(fn->CommInfo))(A, &my_id, &num_procs);
   norms = ((g->logging > 0 || g->print_level > 0) && g->norms) ? g->norms : NULL;
   if (!norms && g->print_level > 1) { norms = own_norms = (double *) calloc((size_t) g->max_iter + 1, sizeof(double)); }

   {
      double *db, *dx;
      if (krylov_operands(pb, px, &db, &dx) != 0) { free(own_norms); return hypre_error_flag; }
      status = db ? b200_gmres_solve(glue_ctx(), m->A, amg, &prm, db, dx, &res, norms)
                  : b200_gmres_solve_host(glue_ctx(), m->A, amg, &prm,
                                          hypre_VectorData(hypre_ParVectorLocalVector(pb)) + m->first_row,
                                          hypre_VectorData(hypre_ParVectorLocalVector(px)) + m->first_row, &res, norms);
   }
   if (status != B200_OK) { free(own_norms); GLUE_FAIL("device COGMRES failed"); return hypre_error_flag; }
   if (verbose() && g_rank == 0)
   {
      fprintf(stderr, "hypre_b200: COGMRES(%d) solve on the device: %d iterations, relative residual %.6e\n",
              (int) g->k_dim, res.num_iterations, res.rel_residual_norm);
   }
   if (res.error_flag & 1)
   {
      if (g->logging > 0 || g->print_level > 0) { krylov_nan_message("hypre_COGMRESSolve"); }
      free(own_norms);
      hypre_error(HYPRE_ERROR_GENERIC);
      return hypre_error_flag;
   }
   if (g->print_level > 1 && my_id == 0 && g_rank == 0)
   {
      hypre_b200_krylov_report(2, norms, res.num_iterations, res.b_norm, res.rel_residual_norm, res.real_residual_exit);
   }
   free(own_norms);
   g->converged = res.converged;
   g->num_iterations = res.num_iterations;
   g->rel_residual_norm = res.rel_residual_norm;
   hypre_ParVectorAllZeros(px) = 0;
   if (res.error_flag & 256) { hypre_error(HYPRE_ERROR_CONV); }
   return hypre_error_flag;
}

/* hypre_FlexGMRESSolve (src/krylov/flexgmres.c:282-…) for the ParCSR operator table, with the default modify_pc
 * (hypre_FlexGMRESModifyPCDefault, flexgmres.c: does nothing -- what ij -solver 61 / 60 installs).  A callback
 * that retunes the preconditioner between iterations has no device counterpart and is rejected. */
HYPRE_Int hypre_FlexGMRESSolve(void *fgmres_vdata, void *A, void *b, void *x)
{
   static pcg_solve_fn next = NULL;
   hypre_FlexGMRESData      *g  = (hypre_FlexGMRESData *) fgmres_vdata;
   hypre_FlexGMRESFunctions *fn = g->functions;
   hypre_ParCSRMatrix       *pA = (hypre_ParCSRMatrix *) A;
   hypre_ParVector          *pb = (hypre_ParVector *) b, *px = (hypre_ParVector *) x;
   mirror_t *m;
   b200_amg *amg = NULL;
   b200_gmres_params prm;
   b200_krylov_result res;
   HYPRE_Int my_id, num_procs;
   double *norms, *own_norms = NULL;
   int status, diag_scale = 0;

   if (!next) { next = (pcg_solve_fn) next_sym("hypre_FlexGMRESSolve"); }
   if (!g_enabled) { return next(fgmres_vdata, A, b, x); }
   if (fn->Matvec != hypre_ParKrylovMatvec) { return next(fgmres_vdata, A, b, x); }

   if (hypre_ParVectorNumVectors(pb) > 1 ||
       (fn->modify_pc != NULL && fn->modify_pc != hypre_FlexGMRESModifyPCDefault))
   {
      GLUE_FAIL("FlexGMRES option outside the device path (multivector / a modify_pc callback)");
      return hypre_error_flag;
   }
   m = krylov_precond_mirror((HYPRE_Int (*)(void *, void *, void *, void *)) fn->precond, g->precond_data, pA, &amg, &diag_scale);
   if (!m) { return hypre_error_flag; }

   memset(&prm, 0, sizeof(prm));
   prm.tol = g->tol; prm.a_tol = g->a_tol; prm.max_iter = g->max_iter; prm.min_iter = g->min_iter;
   prm.k_dim = g->k_dim; prm.diag_scale = diag_scale; prm.flex = 1; prm.cf_tol = g->cf_tol;
   g->converged = 0;
   (*(fn->CommInfo))(A, &my_id, &num_procs);
   norms = ((g->logging > 0 || g->print_level > 0) && g->norms) ? g->norms : NULL;
   if (!norms && g->print_level > 1) { norms = own_norms = (double *) calloc((size_t) g->max_iter + 1, sizeof(double)); }

   {
      double *db, *dx;
      if (krylov_operands(pb, px, &db, &dx) != 0) { free(own_norms); return hypre_error_flag; }
      status = db ? b200_gmres_solve(glue_ctx(), m->A, amg, &prm, db, dx, &res, norms)
                  : b200_gmres_solve_host(glue_ctx(), m->A, amg, &prm,
                                          hypre_VectorData(hypre_ParVectorLocalVector(pb)) + m->first_row,
                                          hypre_VectorData(hypre_ParVectorLocalVector(px)) + m->first_row, &res, norms);
   }
   if (status != B200_OK) { free(own_norms); GLUE_FAIL("device FlexGMRES failed"); return hypre_error_flag; }
   if (verbose() && g_rank == 0)
   {
      fprintf(stderr, "hypre_b200: FlexGMRES(%d) solve on the device: %d iterations, relative residual %.6e\n",
              (int) g->k_dim, res.num_iterations, res.rel_residual_norm);
   }
   if (res.error_flag & 1)
   {
      if (g->logging > 0 || g->print_level > 0) { krylov_nan_message("hypre_FlexGMRESSolve"); }
      free(own_norms);
      hypre_error(HYPRE_ERROR_GENERIC);
      return hypre_error_flag;
   }
   if (g->print_level > 1 && my_id == 0 && g_rank == 0)
   {
      hypre_b200_krylov_report(0, norms, res.num_iterations, res.b_norm, res.rel_residual_norm, res.real_residual_exit);
   }
   free(own_norms);
   g->converged = res.converged;
   g->num_iterations = res.num_iterations;
   g->rel_residual_norm = res.rel_residual_norm;
   hypre_ParVectorAllZeros(px) = 0;
   if (res.error_flag & 256) { hypre_error(HYPRE_ERROR_CONV); }
   return hypre_error_flag;
}

/* hypre_LGMRESSolve (src/krylov/lgmres.c:337-...) for the ParCSR operator table: GMRES augmented with the last
 * aug_dim error approximations (ij -solver 51 / 50). */
HYPRE_Int hypre_LGMRESSolve(void *lgmres_vdata, void *A, void *b, void *x)
{
   static pcg_solve_fn next = NULL;
   hypre_LGMRESData      *g  = (hypre_LGMRESData *) lgmres_vdata;
   hypre_LGMRESFunctions *fn = g->functions;
   hypre_ParCSRMatrix    *pA = (hypre_ParCSRMatrix *) A;
   hypre_ParVector       *pb = (hypre_ParVector *) b, *px = (hypre_ParVector *) x;
   mirror_t *m;
   b200_amg *amg = NULL;
   b200_lgmres_params prm;
   b200_krylov_result res;
   HYPRE_Int my_id, num_procs;
   double *norms, *own_norms = NULL;
   int status, diag_scale = 0;

   if (!next) { next = (pcg_solve_fn) next_sym("hypre_LGMRESSolve"); }
   if (!g_enabled) { return next(lgmres_vdata, A, b, x); }
   if (fn->Matvec != hypre_ParKrylovMatvec) { return next(lgmres_vdata, A, b, x); }

   if (g->k_dim <= g->aug_dim || hypre_ParVectorNumVectors(pb) > 1)
   {
      GLUE_FAIL("LGMRES option outside the device path (aug_dim >= k_dim / multivector)");
      return hypre_error_flag;
   }
   m = krylov_precond_mirror((HYPRE_Int (*)(void *, void *, void *, void *)) fn->precond, g->precond_data, pA, &amg, &diag_scale);
   if (!m) { return hypre_error_flag; }

   memset(&prm, 0, sizeof(prm));
   prm.tol = g->tol; prm.a_tol = g->a_tol; prm.max_iter = g->max_iter; prm.min_iter = g->min_iter;
   prm.k_dim = g->k_dim; prm.aug_dim = g->aug_dim; prm.approx_constant = g->approx_constant; prm.diag_scale = diag_scale;
   prm.cf_tol = g->cf_tol;
   g->converged = 0;
   (*(fn->CommInfo))(A, &my_id, &num_procs);
   norms = ((g->logging > 0 || g->print_level > 0) && g->norms) ? g->norms : NULL;
   if (!norms && g->print_level > 1) { norms = own_norms = (double *) calloc((size_t) g->max_iter + 1, sizeof(double)); }
   {
      double *db, *dx;
      if (krylov_operands(pb, px, &db, &dx) != 0) { free(own_norms); return hypre_error_flag; }
      status = db ? b200_lgmres_solve(glue_ctx(), m->A, amg, &prm, db, dx, &res, norms)
                  : b200_lgmres_solve_host(glue_ctx(), m->A, amg, &prm,
                                           hypre_VectorData(hypre_ParVectorLocalVector(pb)) + m->first_row,
                                           hypre_VectorData(hypre_ParVectorLocalVector(px)) + m->first_row, &res, norms);
   }
   if (status != B200_OK) { free(own_norms); GLUE_FAIL("device LGMRES failed"); return hypre_error_flag; }
   if (verbose() && g_rank == 0)
   {
      fprintf(stderr, "hypre_b200: LGMRES(%d,%d) solve on the device: %d iterations, relative residual %.6e\n",
              (int) g->k_dim, (int) g->aug_dim, res.num_iterations, res.rel_residual_norm);
   }
   if (res.error_flag & 1)
   {
      if (g->logging > 0 || g->print_level > 0) { krylov_nan_message("hypre_LGMRESSolve"); }
      free(own_norms);
      hypre_error(HYPRE_ERROR_GENERIC);
      return hypre_error_flag;
   }
   if (g->print_level > 1 && my_id == 0 && g_rank == 0)
   {
      hypre_b200_krylov_report(0, norms, res.num_iterations, res.b_norm, res.rel_residual_norm, res.real_residual_exit);
   }
   free(own_norms);
   g->converged = res.converged;
   g->num_iterations = res.num_iterations;
   g->rel_residual_norm = res.rel_residual_norm;
   hypre_ParVectorAllZeros(px) = 0;
   if (res.error_flag & 256) { hypre_error(HYPRE_ERROR_CONV); }
   return hypre_error_flag;
}

/* hypre_BiCGSTABSolve (src/krylov/bicgstab.c:226-640) for the ParCSR operator table. */
HYPRE_Int hypre_BiCGSTABSolve(void *bicgstab_vdata, void *A, void *b, void *x)
{
   static pcg_solve_fn next = NULL;
   hypre_BiCGSTABData      *g  = (hypre_BiCGSTABData *) bicgstab_vdata;
   hypre_BiCGSTABFunctions *fn = g->functions;
   hypre_ParCSRMatrix      *pA = (hypre_ParCSRMatrix *) A;
   hypre_ParVector         *pb = (hypre_ParVector *) b, *px = (hypre_ParVector *) x;
   mirror_t *m;
   b200_amg *amg = NULL;
   b200_bicgstab_params prm;
   b200_krylov_result res;
   HYPRE_Int my_id, num_procs;
   double *norms, *own_norms = NULL;
   int status, diag_scale = 0;

   if (!next) { next = (pcg_solve_fn) next_sym("hypre_BiCGSTABSolve"); }
   if (!g_enabled) { return next(bicgstab_vdata, A, b, x); }
   if (fn->Matvec != hypre_ParKrylovMatvec) { return next(bicgstab_vdata, A, b, x); }

   if ((g->hybrid != 0 && g->hybrid != -1) || hypre_ParVectorNumVectors(pb) > 1)
   {
      GLUE_FAIL("BiCGSTAB option outside the device path (multivector)");
      return hypre_error_flag;
   }
   m = krylov_precond_mirror((HYPRE_Int (*)(void *, void *, void *, void *)) fn->precond, g->precond_data, pA, &amg, &diag_scale);
   if (!m) { return hypre_error_flag; }

   memset(&prm, 0, sizeof(prm));
   prm.tol = g->tol; prm.a_tol = g->a_tol; prm.max_iter = g->max_iter; prm.min_iter = g->min_iter;
   prm.diag_scale = diag_scale; prm.stop_crit = g->stop_crit ? 1 : 0; prm.cf_tol = g->cf_tol;
   g->converged = 0;
   (*(fn->CommInfo))(A, &my_id, &num_procs);
   norms = ((g->logging > 0 || g->print_level > 0) && g->norms) ? g->norms : NULL;
   if (!norms && g->print_level > 0) { norms = own_norms = (double *) calloc((size_t) g->max_iter + 1, sizeof(double)); }

   {
      double *db, *dx;
      if (krylov_operands(pb, px, &db, &dx) != 0) { free(own_norms); return hypre_error_flag; }
      status = db ? b200_bicgstab_solve(glue_ctx(), m->A, amg, &prm, db, dx, &res, norms)
                  : b200_bicgstab_solve_host(glue_ctx(), m->A, amg, &prm,
                                             hypre_VectorData(hypre_ParVectorLocalVector(pb)) + m->first_row,
                                             hypre_VectorData(hypre_ParVectorLocalVector(px)) + m->first_row, &res, norms);
   }
   if (status != B200_OK) { free(own_norms); GLUE_FAIL("device BiCGSTAB failed"); return hypre_error_flag; }
   if (verbose() && g_rank == 0)
   {
      fprintf(stderr, "hypre_b200: BiCGSTAB solve on the device: %d iterations, relative residual %.6e\n",
              res.num_iterations, res.rel_residual_norm);
   }
   if (res.error_flag & 1)
   {
      if (g->logging > 0 || g->print_level > 0) { krylov_nan_message("hypre_BiCGSTABSolve"); }
      free(own_norms);
      hypre_error(HYPRE_ERROR_GENERIC);
      return hypre_error_flag;
   }
   if (g->print_level > 0 && my_id == 0 && g_rank == 0)
   {
      hypre_b200_krylov_report(1, norms, res.num_iterations, res.b_norm, res.rel_residual_norm, res.real_residual_exit);
   }
   free(own_norms);
   g->converged = res.converged;
   g->num_iterations = res.num_iterations;
   /* the reference leaves rel_residual_norm untouched when b == 0 and the solve ends before the loop */
   if (res.b_norm > 0.0 || res.num_iterations > 0 || (res.error_flag & 256)) { g->rel_residual_norm = res.rel_residual_norm; }
   hypre_ParVectorAllZeros(px) = 0;
   if (res.error_flag & 2) { hypre_error_w_msg(HYPRE_ERROR_GENERIC, "BiCGSTAB broke down!! divide by near zero\n"); }
   if ((res.error_flag & 256) && g->hybrid != -1) { hypre_error(HYPRE_ERROR_CONV); }
   return hypre_error_flag;
}

/* The closing statistics of hypre_BoomerAMGSolve (src/parcsr_ls/par_amg_solve.c:338-416) and its residual
 * table (:257-264, :327-331) from the history the device solve returned.  hist[k] = ||r_k||_2, k = 0..cycles;
 * rhs_norm = ||f||_2 (0 when it vanishes).  Host arithmetic and printing only; also reachable from the tests
 * (hypre_b200_amg_report) so that the text can be compared with the reference's without a GPU. */
static void amg_solve_report(hypre_ParAMGData *amg_data, hypre_ParCSRMatrix *A, int cycles, const double *hist,
                             double rhs_norm, int not_converged, int table)
{
   const HYPRE_Int nl = hypre_ParAMGDataNumLevels(amg_data);
   hypre_ParCSRMatrix **A_array = hypre_ParAMGDataAArray(amg_data);
   HYPRE_Int *ngs = hypre_ParAMGDataNumGridSweeps(amg_data);
   int sweeps_in[4], k, l;
   long long *sweeps = (long long *) calloc((size_t) nl, sizeof(long long));
   double vars = 0.0, coeffs = 0.0, ops = 0.0, vars0, coeffs0, factor = 1.0;

   if (table)
   {
      for (k = 1; k <= cycles; k++)
      {
         const double f = hist[k - 1] != 0.0 ? hist[k] / hist[k - 1] : hist[k];
         hypre_printf("    Cycle %2d   %e    %f     %e \n", k, hist[k], f, rhs_norm != 0.0 ? hist[k] / rhs_norm : hist[k]);
      }
   }
   for (k = 0; k < 4; k++) { sweeps_in[k] = (int) ngs[k]; }
   b200_cycle_sweep_counts((int) nl, (int) hypre_ParAMGDataCycleType(amg_data), (int) hypre_ParAMGDataFCycle(amg_data),
                           sweeps_in, sweeps);
   vars0 = (double) hypre_ParCSRMatrixGlobalNumRows(A);
   coeffs0 = (double) hypre_ParCSRMatrixDNumNonzeros(A);
   for (l = 0; l < nl; l++)
   {
      /* statistics use the integer count on the coarse levels, the cycle count the double one (as the reference) */
      vars   += l == 0 ? vars0 : (double) hypre_ParCSRMatrixGlobalNumRows(A_array[l]);
      coeffs += l == 0 ? coeffs0 : (double) hypre_ParCSRMatrixNumNonzeros(A_array[l]);
      ops    += (double) sweeps[l] * (double) hypre_ParCSRMatrixDNumNonzeros(A_array[l]);
   }
   free(sweeps);
   if (cycles > 0 && hist[0] != 0.0) { factor = pow(hist[cycles] / hist[0], 1.0 / (double) cycles); }
   if (not_converged)
   {
      hypre_printf("\n\n==============================================");
      hypre_printf("\n NOTE: Convergence tolerance was not achieved\n");
      hypre_printf("      within the allowed %d V-cycles\n", (int) hypre_ParAMGDataMaxIter(amg_data));
      hypre_printf("==============================================");
   }
   hypre_printf("\n\n Average Convergence Factor = %f", factor);
   hypre_printf("\n\n     Complexity:    grid = %f\n", vars0 != 0.0 ? vars / vars0 : 0.0);
   hypre_printf("                operator = %f\n", coeffs0 != 0.0 ? coeffs / coeffs0 : 0.0);
   hypre_printf("                   cycle = %f\n\n\n\n", coeffs0 != 0.0 ? ops / coeffs0 : 0.0);
}

/* test hook: print the table and the statistics for a given history (CPU test of the text) */
void hypre_b200_amg_report(void *amg_vdata, int cycles, const double *hist, double rhs_norm, int not_converged)
{
   hypre_ParAMGData *amg_data = (hypre_ParAMGData *) amg_vdata;
   amg_solve_report(amg_data, hypre_ParAMGDataAArray(amg_data)[0], cycles, hist, rhs_norm, not_converged, 1);
   fflush(stdout);
}

/* hypre_BoomerAMGSolve (src/parcsr_ls/par_amg_solve.c:21-424): the stand-alone AMG iteration of `ij -solver 0`,
 * and the preconditioner call of every Krylov solver of the reference that is not interposed here (GMRES,
 * BiCGSTAB, ...: one device cycle per application, vectors copied in and out). */
HYPRE_Int hypre_BoomerAMGSolve(void *amg_vdata, hypre_ParCSRMatrix *A, hypre_ParVector *f, hypre_ParVector *u)
{
   static amg_setup_fn next = NULL;      /* same signature */
   hypre_ParAMGData *amg_data = (hypre_ParAMGData *) amg_vdata;
   mirror_t *m;
   b200_ctx *ctx;
   double *db = NULL, *dx = NULL, *hist = NULL, *fh, *uh;
   double tol, rel = 1.0, rhs_norm = 0.0;
   HYPRE_Int my_id = 0, print_level, logging, min_iter, max_iter;
   int n_local = 0, cycles = 0, status, track;

   if (!next) { next = (amg_setup_fn) next_sym("hypre_BoomerAMGSolve"); }
   if (!g_enabled) { return next(amg_vdata, A, f, u); }

   print_level = hypre_ParAMGDataPrintLevel(amg_data);
   logging     = hypre_ParAMGDataLogging(amg_data);
   tol         = hypre_ParAMGDataTol(amg_data);
   min_iter    = hypre_ParAMGDataMinIter(amg_data);
   max_iter    = hypre_ParAMGDataMaxIter(amg_data);
   if (hypre_ParVectorNumVectors(f) != hypre_ParVectorNumVectors(u))
   {
      hypre_error_w_msg(HYPRE_ERROR_GENERIC, "Error: num_vectors for RHS and LHS do not match!\n");
      return hypre_error_flag;
   }
   if (hypre_ParVectorNumVectors(f) > 1 || hypre_ParAMGDataConvergeType(amg_data) != 0 || logging > 1)
   {
      GLUE_FAIL("BoomerAMG solve option outside the device path (multivector / converge_type / logging > 1)");
      return hypre_error_flag;
   }
   if (hypre_ParAMGDataAArray(amg_data) == NULL || hypre_ParAMGDataAArray(amg_data)[0] != A)
   {
      GLUE_FAIL("BoomerAMG was not set up for this matrix");
      return hypre_error_flag;
   }
   m = mirror_for_amg(amg_data);
   if (!m) { GLUE_FAIL("hierarchy upload failed"); return hypre_error_flag; }
   ctx = glue_ctx();
   if (g_nranks > 1 && !g_native && tol <= 0.0 && max_iter == 1)
   {
      /* processes holding replicated hypre objects each get their slice of u back: fine for a solve, useless
       * to a host Krylov solver that wants the whole preconditioned vector */
      GLUE_FAIL("BoomerAMG as preconditioner of a host Krylov solver needs the native multi-rank mode");
      return hypre_error_flag;
   }
   hypre_MPI_Comm_rank(hypre_ParCSRMatrixComm(A), &my_id);
   if (g_rank != 0) { my_id = 1; }                    /* replicated processes: only process 0 prints */

   if (b200_parcsr_info(m->A, &n_local, NULL, NULL, NULL, NULL) != B200_OK)
   {
      GLUE_FAIL("device AMG solve failed"); return hypre_error_flag;
   }
   /* this rank's slice of the host vectors (offset 0 unless the processes hold replicated objects) */
   fh = hypre_VectorData(hypre_ParVectorLocalVector(f)) + m->first_row;
   uh = hypre_VectorData(hypre_ParVectorLocalVector(u)) + m->first_row;
   track = (print_level > 1 || tol > 0.0);
   if (track) { hist = (double *) calloc((size_t) (max_iter > 0 ? max_iter : 0) + 2, sizeof(double)); }
   status = b200_dev_alloc(ctx, n_local > 0 ? n_local : 1, &db);
   if (status == B200_OK) { status = b200_dev_alloc(ctx, n_local > 0 ? n_local : 1, &dx); }
   if (status == B200_OK && n_local > 0) { status = b200_h2d(ctx, db, fh, n_local); }
   if (status == B200_OK && n_local > 0) { status = b200_h2d(ctx, dx, uh, n_local); }

   if (status == B200_OK && my_id == 0 && print_level > 1)
   {
      hypre_BoomerAMGWriteSolverParams(amg_data);
      if (tol > 0.0) { hypre_printf("\n\nAMG SOLUTION INFO:\n"); }
   }
   if (status == B200_OK)
   {
      status = b200_amg_solve(m->amg, db, dx, tol, (int) min_iter, (int) max_iter, &cycles, &rel, hist);
   }
   if (status == B200_ERR_NUMERIC)
   {
      if (print_level > 0)
      {
         hypre_printf("\n\nERROR detected by Hypre ...  BEGIN\n");
         hypre_printf("ERROR -- hypre_BoomerAMGSolve: INFs and/or NaNs detected in input.\n");
         hypre_printf("User probably placed non-numerics in supplied A, x_0, or b.\n");
         hypre_printf("ERROR detected by Hypre ...  END\n\n\n");
      }
      hypre_error(HYPRE_ERROR_GENERIC);
   }
   else if (status == B200_OK && n_local > 0) { status = b200_d2h(ctx, uh, dx, n_local); }
   if (db) { b200_dev_free(ctx, db); }
   if (dx) { b200_dev_free(ctx, dx); }
   if (status != B200_OK)
   {
      if (status != B200_ERR_NUMERIC) { GLUE_FAIL("device AMG solve failed"); }
      free(hist);
      return hypre_error_flag;
   }

   hypre_ParAMGDataNumIterations(amg_data) = cycles;
   if (track) { hypre_ParAMGDataRelativeResidualNorm(amg_data) = rel; }
   hypre_ParVectorAllZeros(u) = 0;
   if (cycles == max_iter && tol > 0.0) { hypre_error(HYPRE_ERROR_CONV); }

   if (track)
   {
      /* ||f|| from the last residual and its relative value (rel = ||r|| / ||f||, or ||r|| when f == 0) */
      if (rel != 0.0 && hist[cycles] != rel) { rhs_norm = hist[cycles] / rel; }
      if (tol <= 0.0 && rhs_norm != 0.0)
      {
         /* the reference skips the first residual when no tolerance is set and reports ||f|| instead
          * (par_amg_solve.c:176-192) */
         hist[0] = rhs_norm;
      }
   }
   if (my_id == 0 && print_level > 1)
   {
      hypre_printf("                                            relative\n");
      hypre_printf("               residual        factor       residual\n");
      hypre_printf("               --------        ------       --------\n");
      hypre_printf("    Initial    %e                 %e\n", hist[0], rhs_norm != 0.0 ? hist[0] / rhs_norm : hist[0]);
      amg_solve_report(amg_data, A, cycles, hist, rhs_norm, cycles == max_iter && tol > 0.0, 1);
   }
   if (verbose() && my_id == 0)
   {
      fprintf(stderr, "hypre_b200: BoomerAMG solve on the device: %d cycles, relative residual %.6e\n", cycles, rel);
   }
   free(hist);
   return hypre_error_flag;
}

/* Public SpMV y = alpha*A*x + beta*y (src/parcsr_mv/HYPRE_parcsr_matrix.c), the call
 * `ij -solver -1 -nmv N` times (src/test/ij.c:4262-4296).  The matrix is mirrored on the
 * first call; library-internal hypre_ParCSRMatrixMatvec calls (host setup) are untouched. */
HYPRE_Int HYPRE_ParCSRMatrixMatvec(HYPRE_Complex alpha, HYPRE_ParCSRMatrix A, HYPRE_ParVector x,
                                   HYPRE_Complex beta, HYPRE_ParVector y)
{
   hypre_ParCSRMatrix *pA = (hypre_ParCSRMatrix *) A;
   hypre_ParVector    *px = (hypre_ParVector *) x, *py = (hypre_ParVector *) y;
   mirror_t *m;
   if (!g_enabled)
   {
      typedef HYPRE_Int (*mv_fn)(HYPRE_Complex, HYPRE_ParCSRMatrix, HYPRE_ParVector, HYPRE_Complex, HYPRE_ParVector);
      static mv_fn next = NULL;
      if (!next) { next = (mv_fn) next_sym("HYPRE_ParCSRMatrixMatvec"); }
      return next(alpha, A, x, beta, y);
   }
   if (hypre_ParVectorNumVectors(px) > 1)
   {
      GLUE_FAIL("multivector SpMV is outside the device path");
      return hypre_error_flag;
   }
   m = mirror_for_matrix(pA);
   if (!m) { GLUE_FAIL("matrix upload failed"); return hypre_error_flag; }
   {
      /* device-bound operands (hypre_b200_vector_bind_device): the product runs on them in place */
      double *dx, *dy;
      int st;
      if (krylov_operands(px, py, &dx, &dy) != 0) { return hypre_error_flag; }
      st = dx ? b200_parcsr_matvec(m->A, alpha, dx, beta, dy, dy)
              : b200_parcsr_matvec_host(m->A, alpha, hypre_VectorData(hypre_ParVectorLocalVector(px)) + m->first_col, beta,
                                        hypre_VectorData(hypre_ParVectorLocalVector(py)) + m->first_row,
                                        hypre_VectorData(hypre_ParVectorLocalVector(py)) + m->first_row);
      if (st != B200_OK) { GLUE_FAIL("device SpMV failed"); }
   }
   hypre_ParVectorAllZeros(py) = 0;
   return hypre_error_flag;
}

/* handles for tests / bench (python ctypes) */
/* y = alpha A^T x + beta y for host-resident hypre vectors: HYPRE_ParCSRMatrixMatvecT
 * (src/parcsr_mv/HYPRE_parcsr_matrix.c) -> hypre_ParCSRMatrixMatvecT (par_csr_matvec.c:522-546).  x lives on the
 * row partition, y on the column partition of A. */
HYPRE_Int HYPRE_ParCSRMatrixMatvecT(HYPRE_Complex alpha, HYPRE_ParCSRMatrix A, HYPRE_ParVector x,
                                    HYPRE_Complex beta, HYPRE_ParVector y)
{
   hypre_ParCSRMatrix *pA = (hypre_ParCSRMatrix *) A;
   hypre_ParVector    *px = (hypre_ParVector *) x, *py = (hypre_ParVector *) y;
   mirror_t *m;
   if (!g_enabled)
   {
      typedef HYPRE_Int (*mv_fn)(HYPRE_Complex, HYPRE_ParCSRMatrix, HYPRE_ParVector, HYPRE_Complex, HYPRE_ParVector);
      static mv_fn next = NULL;
      if (!next) { next = (mv_fn) next_sym("HYPRE_ParCSRMatrixMatvecT"); }
      return next(alpha, A, x, beta, y);
   }
   if (hypre_ParVectorNumVectors(px) > 1)
   {
      GLUE_FAIL("multivector SpMV is outside the device path");
      return hypre_error_flag;
   }
   m = mirror_for_matrix(pA);
   if (!m) { GLUE_FAIL("matrix upload failed"); return hypre_error_flag; }
   {
      double *dx, *dy;
      int st;
      if (krylov_operands(px, py, &dx, &dy) != 0) { return hypre_error_flag; }
      st = dx ? b200_parcsr_matvecT(m->A, alpha, dx, beta, dy)
              : b200_parcsr_matvecT_host(m->A, alpha, hypre_VectorData(hypre_ParVectorLocalVector(px)) + m->first_row, beta,
                                         hypre_VectorData(hypre_ParVectorLocalVector(py)) + m->first_col);
      if (st != B200_OK) { GLUE_FAIL("device SpMV^T failed"); }
   }
   hypre_ParVectorAllZeros(py) = 0;
   return hypre_error_flag;
}

b200_amg *hypre_b200_amg_mirror(void *amg_vdata)
{
   mirror_t *m = mirror_for_amg((hypre_ParAMGData *) amg_vdata);
   return m ? m->amg : NULL;
}
b200_parcsr *hypre_b200_matrix_mirror(void *parcsr)
{
   mirror_t *m = mirror_for_matrix((hypre_ParCSRMatrix *) parcsr);
   return m ? m->A : NULL;
}
/* distributed mode: row starts (nranks+1 entries) of `level` of the mirrored hierarchy.  Returns the
 * number of row blocks: nranks for a distributed level, 1 for a replicated level or a single rank. */
int hypre_b200_amg_level_starts(void *amg_vdata, int level, int *out)
{
   mirror_t *m = mirror_find(amg_vdata);
   int r;
   if (!m || !m->starts || level < 0 || level >= m->num_levels || level >= m->repl_level) { return 1; }
   for (r = 0; r <= g_nranks; r++) { out[r] = m->starts[(size_t) level * (g_nranks + 1) + r]; }
   return g_nranks;
}
/* this rank's first row of the level-0 operator mirrored for `key` (amg data or matrix) */
int hypre_b200_first_row(void *key)
{
   mirror_t *m = mirror_find(key);
   return m ? m->first_row : -1;
}
