// pcg.cu -- preconditioned conjugate gradients on the device
// (hypre_PCGSolve, src/krylov/pcg.c:317-1008, with the ParCSR operator table of
// src/parcsr_ls/par_krylov_func.c).  Scalars stay on the device between the
// kernels of one iteration; the host reads {sdotp, <r,r>, <r,Cr>, breakdown flag}
// once per iteration for the reference's convergence / breakdown logic.
#include "ops.cuh"
#include "vecops.cuh"
#include "p2p.cuh"
#include "tail.cuh"
#include <cmath>
#include <cfloat>

int     amg_precond_ptr(b200_amg *a, const double *f, double *u, double **u_result, const FusedDot *fd = nullptr);
int     amg_precond_cycles(const b200_amg *a);

// scalar slots in ctx->d_scalars
enum { S_SDOTP = 8, S_IPROD = 9, S_GAMMA_NEW = 10, S_GAMMA = 11, S_BETA = 12, S_BIPROD = 13, S_EPS = 14, S_TMP = 15,
       S_EPOCH = 20 /* solve counter (exact in a double), tags the records of this solve */ };

// per-iteration record the device writes into mapped pinned host memory (zero-copy): the host
// follows the iteration without any stream synchronisation
enum { PCG_CONVERGED = 1, PCG_ERR_CONV = 2, PCG_BRK_PRE = 4, PCG_BRK_POST = 8 };
struct PcgRecord {
   volatile long long ready;   // (epoch << 32) | iteration, written last
   int    stop, flags;
   double sdotp, i_prod, gamma_new;
};
constexpr int PCG_RING = 8;

// x += alpha p ; r -= alpha s ; red0 = <r,r>   with alpha = gamma / sdotp from device-resident
// scalars; the reference's breakdown guards (pcg.c:581-629) are evaluated here as well so that x and
// r stay untouched when the reference would `break`.
struct FPcgUpdate {
   static constexpr int NRED = 1;
   int skip_break; const double *scal; const double *p; const double *s; double *x; double *r;
   double alpha; bool stop;      // set by init(), once per thread
   const int *iter; const double *p_alt;     // fused-direction mode: p is the buffer EpiMatvecDotDir just wrote
   __device__ void init()
   {
      if (iter != nullptr && (*iter & 1)) { p = p_alt; }
      const double sdotp = scal[S_SDOTP];
      alpha = scal[S_GAMMA] / sdotp;
      stop = (sdotp == 0.0);
      if (!stop)
      {
         if (alpha <= 0.0)               { stop = skip_break < 3; }
         else if (!(alpha >= 4.9406564584124654e-324)) { stop = skip_break < 2; }
         else if (!(alpha >= DBL_MIN))   { stop = skip_break < 1; }
      }
   }
   __device__ void operator()(int i, double *red) const
   {
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

// p = z + beta p   (pcg.c:960-971: ScaleVector(beta,p); Axpy(1,s,p)), beta from the device
struct FPcgDirection {
   static constexpr int NRED = 0;
   const double *scal; const double *z; double *p;
   double beta;
   __device__ void init() { beta = scal[S_BETA]; }
   __device__ void operator()(int i, double *) const { p[i] = beta * p[i] + z[i]; }
};

// z = r ./ d  (HYPRE_ParCSRDiagScale) and red0 = <r,z>
struct FDiagScaleDot {
   static constexpr int NRED = 1; __device__ void init() {}
   const double *d; const double *r; double *z;
   __device__ void operator()(int i, double *red) const
   {
      double zi = r[i] / d[i];
      z[i] = zi;
      red[0] = fma(r[i], zi, red[0]);
   }
};

// The scalar logic of one iteration of hypre_PCGSolve (pcg.c:581-629 alpha guards, :745-843
// convergence test, :845-883 gamma guards, :960 beta) on the device: one thread.  Sets the stop
// word that turns every later kernel into a no-op, publishes the record for the host.
__global__ void pcg_check_kernel(double *scal, int *stop, int *iter, PcgRecord *ring,
                                 int two_norm, int skip_break, int max_iter)
{
   if (*stop != 0) { return; }
   const long long epoch = (long long)scal[S_EPOCH];
   const int i = *iter + 1;
   const double sdotp = scal[S_SDOTP], gamma = scal[S_GAMMA];
   int flags = 0, st = 0;
   double gamma_new = 0.0, i_prod = 0.0;
   bool brk = false;
   if (sdotp == 0.0) { flags |= PCG_ERR_CONV; brk = true; }
   else
   {
      const double alpha = gamma / sdotp;
      if (alpha <= 0.0)                              { flags |= PCG_ERR_CONV; brk = skip_break < 3; }
      else if (!(alpha >= 4.9406564584124654e-324))  { flags |= PCG_ERR_CONV; brk = skip_break < 2; }
      else if (!(alpha >= DBL_MIN))                  { flags |= PCG_ERR_CONV; brk = skip_break < 1; }
   }
   if (brk) { flags |= PCG_BRK_PRE; st = 1; }
   else
   {
      gamma_new = scal[S_GAMMA_NEW];
      i_prod = two_norm ? scal[S_IPROD] : gamma_new;
      if (i_prod / scal[S_BIPROD] < scal[S_EPS]) { flags |= PCG_CONVERGED; st = 1; }
      else
      {
         bool gb = false;
         if (gamma_new <= 0.0)                              { flags |= PCG_ERR_CONV; gb = skip_break < 3; }
         else if (!(gamma_new >= 4.9406564584124654e-324))  { flags |= PCG_ERR_CONV; gb = skip_break < 2; }
         else if (!(gamma_new >= DBL_MIN))                  { flags |= PCG_ERR_CONV; gb = skip_break < 1; }
         if (gb) { flags |= PCG_BRK_POST; st = 1; }
         else
         {
            scal[S_BETA] = gamma_new / gamma;
            scal[S_GAMMA] = gamma_new;
            if (i >= max_iter) { st = 1; }
         }
      }
   }
   *iter = i;
   if (st) { *stop = 1; }
   PcgRecord *rec = ring + (i % PCG_RING);
   rec->stop = st; rec->flags = flags; rec->sdotp = sdotp; rec->i_prod = i_prod; rec->gamma_new = gamma_new;
   __threadfence_system();
   rec->ready = (epoch << 32) | (long long)i;
}

static int dot_to(b200_ctx *c, int n, const double *x, const double *y, int slot)
{
   return vec_dot(c, n, x, y, c->d_scalars + slot);
}

// z = C r.  Returns in *z_out the buffer that holds the result (z or the AMG's level-0 alternate).
// dot_slot >= 0: <r, z> is wanted in that scalar slot; *dot_done tells whether the preconditioner's last launch
// reduced it (the caller runs a dot kernel otherwise).
static int apply_precond(b200_ctx *c, b200_parcsr *A, b200_amg *amg, const b200_pcg_params *prm,
                         const double *r, double *z, double **z_out, int dot_slot = -1, bool *dot_done = nullptr)
{
   const int n = A->n_rows;
   *z_out = z;
   bool done_local = false;
   FusedDot fd{r, dot_slot >= 0 ? c->d_scalars + dot_slot : nullptr, dot_done ? dot_done : &done_local};
   if (dot_done) { *dot_done = false; }
   if (amg) { return amg_precond_ptr(amg, r, z, z_out, dot_slot >= 0 ? &fd : nullptr); }
   if (prm->diag_scale)
   {
      B200_TRY(parcsr_ensure_L(A));
      FDiagScaleDot f{A->d_diagonal, r, z};
      if (dot_slot >= 0) { *fd.done = true; }
      return run_ewise(c, n, f, c->d_scalars + (dot_slot >= 0 ? dot_slot : S_TMP));
   }
   return b200_vec_copy(c, n, r, z);   // identity: hypre_ParKrylovIdentity
}

static const int g_pcg_fuse = []() { const char *e = getenv("B200_PCG_FUSE"); return e ? atoi(e) : 1; }();

// ---- one iteration, enqueued on the compute stream without any host synchronisation
struct PcgIter {
   b200_ctx *c; b200_parcsr *A; b200_amg *amg; const b200_pcg_params *prm;
   int n; double *p, *s, *r, *x; long long epoch;
   // fused mode (single rank or peer-memory halo): the direction update lives inside s = A p (two p buffers
   // alternate with the parity of the device's iteration counter), <r, z> inside the preconditioner's last launch
   bool fused; double *p2, *zbuf; double *z;     // z: where z = C r of the PREVIOUS step lives
};

// One iteration in fused form: 3 launches around the preconditioner instead of 5 (no stand-alone dot, no direction
// pass): [s = A (z + beta p), p' = z + beta p, <s,p'>] [x += alpha p', r -= alpha s, <r,r>] [z = C r with <r,z>] [check]
static int pcg_enqueue_iteration_fused(const PcgIter &it)
{
   b200_ctx *c = it.c;
   B200_TRY(op_matvec_dot_direction(it.A, c->d_scalars + S_BETA, c->d_iter, it.z, it.p, it.p2, it.s, c->d_scalars + S_SDOTP));
   B200_TRY(scalars_allreduce(c, c->d_scalars + S_SDOTP, 1));
   {
      // even count of finished iterations: the new p is p2 (EpiMatvecDotDir::init)
      FPcgUpdate f{it.prm->skip_break, c->d_scalars, it.p2, it.s, it.x, it.r, 0.0, false, c->d_iter, it.p};
      B200_TRY(run_ewise(c, it.n, f, c->d_scalars + S_IPROD));
   }
   double *z = nullptr;
   bool dot_done = false;
   B200_TRY(apply_precond(c, it.A, it.amg, it.prm, it.r, it.zbuf, &z, S_GAMMA_NEW, &dot_done));
   if (z != it.z)
   {
      b200_set_error("PCG: the preconditioner's result moved between applications");
      return B200_ERR_ARG;
   }
   if (!dot_done) { B200_TRY(vec_dot(c, it.n, it.r, z, c->d_scalars + S_GAMMA_NEW)); }
   B200_TRY(scalars_allreduce(c, c->d_scalars + S_IPROD, 2));
   LAUNCH(c, pcg_check_kernel, 1, 1, 0, c->d_scalars, c->d_stop, c->d_iter, (PcgRecord *)c->d_ring,
          it.prm->two_norm, it.prm->skip_break, it.prm->max_iter);
   CHECK_LAUNCH();
   return B200_OK;
}

static int pcg_enqueue_iteration(const PcgIter &it)
{
   if (it.fused) { return pcg_enqueue_iteration_fused(it); }
   b200_ctx *c = it.c;
   double *z = nullptr;
   // s = A p ; sdotp = <s,p>
   B200_TRY(op_matvec_dot(it.A, it.p, it.s, c->d_scalars + S_SDOTP));
   B200_TRY(scalars_allreduce(c, c->d_scalars + S_SDOTP, 1));
   // x += alpha p ; r -= alpha s ; <r,r>
   {
      FPcgUpdate f{it.prm->skip_break, c->d_scalars, it.p, it.s, it.x, it.r, 0.0, false, nullptr, nullptr};
      B200_TRY(run_ewise(c, it.n, f, c->d_scalars + S_IPROD));
   }
   // s = C r ; gamma = <r,s> (reduced inside the preconditioner's last launch where it can be)
   bool dot_done = false;
   if (g_pcg_fuse & 1) { B200_TRY(apply_precond(c, it.A, it.amg, it.prm, it.r, it.s, &z, S_GAMMA_NEW, &dot_done)); }
   else
   {
      B200_TRY(apply_precond(c, it.A, it.amg, it.prm, it.r, it.s, &z));
      if (!it.amg && it.prm->diag_scale)
      {
         CUDA_TRY(cudaMemcpyAsync(c->d_scalars + S_GAMMA_NEW, c->d_scalars + S_TMP, sizeof(double),
                                  cudaMemcpyDeviceToDevice, c->stream));
         dot_done = true;
      }
   }
   if (!dot_done) { B200_TRY(dot_to(c, it.n, it.r, z, S_GAMMA_NEW)); }
   B200_TRY(scalars_allreduce(c, c->d_scalars + S_IPROD, 2));
   // scalar logic + record for the host
   LAUNCH(c, pcg_check_kernel, 1, 1, 0, c->d_scalars, c->d_stop, c->d_iter, (PcgRecord *)c->d_ring,
          it.prm->two_norm, it.prm->skip_break, it.prm->max_iter);
   CHECK_LAUNCH();
   // p = s + beta p
   {
      FPcgDirection f{c->d_scalars, z, it.p, 0.0};
      B200_TRY(run_ewise(c, it.n, f, nullptr));
   }
   return B200_OK;
}

// The iteration is identical from one trip to the next (all scalars live on the device), so it is
// captured once into a CUDA graph and replayed: one graph launch instead of ~75 kernel launches.
static bool pcg_graph_enabled(const b200_ctx *c)
{
   static int env = -1;
   if (env < 0) { const char *e = getenv("B200_GRAPH"); env = e ? atoi(e) : 1; }
   // the NCCL halo path forks onto the comm stream; keep it on plain launches
   return env != 0 && !c->prof && (c->nranks == 1 || c->use_p2p);
}

static int pcg_launch_iteration(const PcgIter &it, int i)
{
   b200_ctx *c = it.c;
   // iteration 1 runs on plain launches: it triggers every lazy allocation (transposes, workspaces)
   if (!pcg_graph_enabled(c) || i < 2) { return pcg_enqueue_iteration(it); }
   const void *key[8] = {it.A, it.amg, it.x, it.p, it.s, it.r,
                         (const void *)(size_t)(it.prm->two_norm * 64 + it.prm->skip_break * 8 + it.prm->diag_scale +
                                                (it.fused ? 128 : 0) + (it.amg ? 4096 * amg_precond_cycles(it.amg) : 0)),
                         (const void *)(size_t)((long long)it.prm->max_iter + (c->generation << 20))};
   if (c->pcg_graph == nullptr || memcmp(key, c->pcg_graph_key, sizeof(key)) != 0)
   {
      if (c->pcg_graph) { cudaGraphExecDestroy(c->pcg_graph); c->pcg_graph = nullptr; }
      cudaGraph_t g = nullptr;
      const long long l0 = c->launches;
      CUDA_TRY(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
      c->capturing = true;
      int st = pcg_enqueue_iteration(it);
      c->capturing = false;
      cudaError_t e = cudaStreamEndCapture(c->stream, &g);
      c->pcg_graph_launches = c->launches - l0;
      c->launches = l0;
      if (st != B200_OK) { if (g) { cudaGraphDestroy(g); } return st; }
      CUDA_TRY(e);
      CUDA_TRY(cudaGraphInstantiate(&c->pcg_graph, g, 0));
      CUDA_TRY(cudaGraphDestroy(g));
      memcpy(c->pcg_graph_key, key, sizeof(key));
   }
   CUDA_TRY(cudaGraphLaunch(c->pcg_graph, c->stream));
   c->launches += c->pcg_graph_launches;
   return B200_OK;
}

// wait (host spin on mapped pinned memory) until the device published iteration i of this solve
static int pcg_wait_record(b200_ctx *c, long long epoch, int i, PcgRecord *out)
{
   volatile PcgRecord *rec = (volatile PcgRecord *)c->h_ring + (i % PCG_RING);
   const long long want = (epoch << 32) | (long long)i;
   unsigned long long spins = 0;
   while (rec->ready != want)
   {
      if ((++spins & 0x3fff) == 0)
      {
         cudaError_t q = cudaStreamQuery(c->stream);
         if (q != cudaSuccess && q != cudaErrorNotReady) { CUDA_TRY(q); }
         if (q == cudaSuccess && rec->ready != want)
         {
            // stream drained without publishing: the stop word was already set (cannot happen in a
            // healthy run) -- report instead of spinning forever
            b200_set_error("PCG iteration %d was never published by the device", i);
            return B200_ERR_CUDA;
         }
      }
   }
   out->stop = rec->stop; out->flags = rec->flags; out->sdotp = rec->sdotp; out->i_prod = rec->i_prod;
   out->gamma_new = rec->gamma_new;
   return B200_OK;
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

   double *p = nullptr, *s = nullptr, *r = nullptr, *p2 = nullptr, *zbuf = nullptr;
   B200_TRY(ctx_workspace(c, WS_PCG_P, n, &p));
   B200_TRY(ctx_workspace(c, WS_PCG_S, n, &s));
   B200_TRY(ctx_workspace(c, WS_PCG_R, n, &r));
   // B200_PCG_FUSE: bit 0 = <r,z> inside the preconditioner's last launch (default), bit 1 = direction update inside
   // s = A p (measured slower on the pattern kernel, which is bound by gathers, not by bytes: off by default)
   const bool fused = (g_pcg_fuse & 2) != 0 && op_can_fuse_gather(A);
   if (fused)
   {
      B200_TRY(ctx_workspace(c, WS_PCG_P2, n, &p2));
      B200_TRY(ctx_workspace(c, WS_PCG_Z, n, &zbuf));
   }
   int status = B200_OK;
   int i = 0, queued = 0;
   double bi_prod = 0.0, i_prod = 0.0, i_prod_0 = 0.0, gamma = 0.0, eps = 0.0;
   double *z = nullptr;
   const long long epoch = ++c->pcg_epoch;
   PcgIter it{c, A, amg, prm, n, p, s, r, x, epoch, fused, p2, zbuf, nullptr};
   constexpr int LOOKAHEAD = 1;     // iterations queued beyond the one the host is waiting for

#define PCG_TRY(expr) do { status = (expr); if (status != B200_OK) goto done; } while (0)

   CUDA_TRY(cudaMemsetAsync(c->d_stop, 0, sizeof(int), c->stream));
   CUDA_TRY(cudaMemsetAsync(c->d_iter, 0, sizeof(int), c->stream));
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
   if (fused)
   {
      // z = C r stays where the preconditioner leaves it; the first fused SpMV forms p = z + 0 * 0
      PCG_TRY(apply_precond(c, A, amg, prm, r, zbuf, &z));
      it.z = z;
      PCG_TRY(vec_fill(c, n, 0.0, p));
      CUDA_TRY(cudaMemsetAsync(c->d_scalars + S_BETA, 0, sizeof(double), c->stream));
      PCG_TRY(dot_to(c, n, r, z, S_GAMMA));
   }
   else
   {
      PCG_TRY(apply_precond(c, A, amg, prm, r, p, &z));
      if (z != p) { PCG_TRY(b200_vec_copy(c, n, z, p)); }
      PCG_TRY(dot_to(c, n, r, p, S_GAMMA));
   }
   PCG_TRY(dot_to(c, n, r, r, S_IPROD));
   // S_IPROD (9) and S_GAMMA (11) are not adjacent: reduce the span 9..11 (S_GAMMA_NEW is scratch here)
   PCG_TRY(scalars_allreduce(c, c->d_scalars + S_IPROD, 3));
   PCG_TRY(scalars_to_host(c, S_IPROD, 3));
   gamma = h[S_GAMMA];
   if (gamma != 0.0 && !(gamma / gamma == gamma / gamma)) { res->error_flag = 1; goto done; }
   i_prod_0 = two_norm ? h[S_IPROD] : gamma;
   if (norms_h) { norms_h[0] = sqrt(i_prod_0); }
   // device-resident constants of the stopping rule
   h[S_BIPROD] = bi_prod; h[S_EPS] = eps; h[S_EPOCH] = (double)epoch;
   CUDA_TRY(cudaMemcpyAsync(c->d_scalars + S_BIPROD, h + S_BIPROD, 2 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
   CUDA_TRY(cudaMemcpyAsync(c->d_scalars + S_EPOCH, h + S_EPOCH, sizeof(double), cudaMemcpyHostToDevice, c->stream));

   while (i < prm->max_iter)
   {
      // keep the device fed: queue up to LOOKAHEAD iterations beyond the one we wait for
      while (queued < prm->max_iter && queued <= i + LOOKAHEAD) { PCG_TRY(pcg_launch_iteration(it, ++queued)); }
      PcgRecord rec;
      PCG_TRY(pcg_wait_record(c, epoch, i + 1, &rec));
      i++;
      if (rec.flags & PCG_ERR_CONV) { res->error_flag |= 256; }
      if (rec.flags & PCG_BRK_PRE) { if (i == 1) { i_prod = i_prod_0; } break; }
      gamma = rec.gamma_new;
      i_prod = rec.i_prod;
      if (norms_h) { norms_h[i] = sqrt(i_prod); }
      if (rel_norms_h) { rel_norms_h[i] = bi_prod ? sqrt(i_prod / bi_prod) : 0.0; }
      if (rec.flags & PCG_CONVERGED) { res->converged = 1; break; }
      if (rec.flags & PCG_BRK_POST) { if (i == 1) { i_prod = i_prod_0; } break; }
   }
   if (i >= prm->max_iter && (i_prod / bi_prod) >= eps && eps > 0) { res->error_flag |= 256; }
   res->num_iterations = i;
   res->rel_residual_norm = bi_prod > 0.0 ? sqrt(i_prod / bi_prod) : 0.0;
   (void)gamma;

done:
#undef PCG_TRY
   // drain (iterations queued past the stop are no-ops), then re-arm the kernels for other callers
   cudaStreamSynchronize(c->stream);
   cudaMemsetAsync(c->d_stop, 0, sizeof(int), c->stream);
   cudaStreamSynchronize(c->stream);
   if (status == B200_OK && c->use_p2p && c->h_err && *c->h_err) { status = B200_ERR_NCCL; }
   if (status == B200_OK && tail_error(c))
   {
      b200_set_error("a grid barrier of the persistent coarse-level kernel timed out");
      status = B200_ERR_CUDA;
   }
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
