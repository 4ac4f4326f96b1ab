// krylov.cu -- restarted GMRES and BiCGSTAB on the device
// (hypre_GMRESSolve, src/krylov/gmres.c:274-930; hypre_BiCGSTABSolve, src/krylov/bicgstab.c:226-640; ParCSR
// operator table of src/parcsr_ls/par_krylov_func.c).  The loops and their scalar logic live in
// krylov_core.hpp; this file is the device backend: every vector operation is one launch of the streaming
// kernels of vecops.cuh / the fused row kernels of ops.cu, dot products stay on the device until the one
// place per iteration where the host needs them (Hessenberg column, step lengths).
//
// Algorithmic HBM bytes per inner iteration (n rows, besides the SpMV and the preconditioner):
//   GMRES step i:   modified Gram-Schmidt as i fused "subtract + next dot" passes (24 n each, the reference
//                   does i dots + i axpys = 40 n each) + first dot 16 n + normalisation 16 n
//   BiCGSTAB:       32 n (x, r half step) + 16 n + 16 n (dots) + 56 n (full step with both dots) + 32 n (p)
#include "ops.cuh"
#include "vecops.cuh"
#include "p2p.cuh"
#include "krylov_core.hpp"

int amg_precond_ptr(b200_amg *a, const double *f, double *u, double **u_result, const FusedDot *fd = nullptr);

// z = r ./ d   (HYPRE_ParCSRDiagScale, src/parcsr_ls/HYPRE_parcsr_pcg.c:267-277)
struct FDiagScale {
   static constexpr int NRED = 0; __device__ void init() {}
   const double *d; const double *r; double *z;
   __device__ void operator()(int i, double *) const { z[i] = r[i] / d[i]; }
};

// one step of modified Gram-Schmidt fused with the next inner product:
//   pi -= h pj ; red0 = <pn, pi>     (pn == pi: the squared norm of the updated vector)
// h = the inner product the previous launch left on the device (already summed over the ranks)
struct FMgsStep {
   static constexpr int NRED = 1;
   const double *h_dev; const double *pj; const double *pn; double *pi;
   double h;
   __device__ void init() { h = *h_dev; }
   __device__ void operator()(int i, double *red) const
   {
      const double v = fma(-h, pj[i], pi[i]);
      pi[i] = v;
      const double o = (pn == pi) ? v : pn[i];
      red[0] = fma(o, v, red[0]);
   }
};

// Classical Gram-Schmidt (COGMRES), four basis vectors per pass:
//   red[q] = <pi, pj[q]>                                   (hypre_ParKrylovMassInnerProd: one all-reduce for all)
struct FMassDot4 {
   static constexpr int NRED = 4; __device__ void init() {}
   int cnt; const double *pi; const double *pj[4];
   __device__ void operator()(int i, double *red) const
   {
      const double v = pi[i];
#pragma unroll
      for (int q = 0; q < 4; q++) { if (q < cnt) { red[q] = fma(v, pj[q][i], red[q]); } }
   }
};
//   pi += sum_q (-h[q]) pj[q]   in the order q = 0..cnt-1 (hypre_SeqVectorMassAxpy, vector_batched.c:236-246),
//   and on the last pass red0 = <pi, pi>
struct FMassAxpy4 {
   static constexpr int NRED = 1;
   int cnt, last; const double *h_dev; const double *pj[4]; double *pi;
   double h[4];
   __device__ void init()
   {
#pragma unroll
      for (int q = 0; q < 4; q++) { h[q] = q < cnt ? h_dev[q] : 0.0; }
   }
   __device__ void operator()(int i, double *red) const
   {
      double v = pi[i];
#pragma unroll
      for (int q = 0; q < 4; q++) { if (q < cnt) { v = fma(-h[q], pj[q][i], v); } }
      pi[i] = v;
      if (last) { red[0] = fma(v, v, red[0]); }
   }
};

// x += a v ; r -= a q
struct FBicgHalf {
   static constexpr int NRED = 0; __device__ void init() {}
   double a; const double *v; const double *q; double *x; double *r;
   __device__ void operator()(int i, double *) const
   {
      x[i] = fma(a, v[i], x[i]);
      r[i] = fma(-a, q[i], r[i]);
   }
};

// x += g v ; r -= g s ; red0 = <r,r> ; red1 = <r0,r>
struct FBicgFull {
   static constexpr int NRED = 2; __device__ void init() {}
   double g; const double *v; const double *s; const double *r0; double *x; double *r;
   __device__ void operator()(int i, double *red) const
   {
      x[i] = fma(g, v[i], x[i]);
      const double ri = fma(-g, s[i], r[i]);
      r[i] = ri;
      red[0] = fma(ri, ri, red[0]);
      red[1] = fma(r0[i], ri, red[1]);
   }
};

// p = f (p - g q) + r
struct FBicgDir {
   static constexpr int NRED = 0; __device__ void init() {}
   double g, f; const double *q; const double *r; double *p;
   __device__ void operator()(int i, double *) const { p[i] = fma(f, fma(-g, q[i], p[i]), r[i]); }
};

// ---------------------------------------------------------------- device backend of krylov_core.hpp
constexpr int AR_CHUNK = 128;     // scalars per all-reduce call (p2p_allreduce takes at most 256)
struct DevKrylovOps {
   b200_ctx *c; b200_parcsr *A; b200_amg *amg; int diag_scale; int n;
   double *d_sc;      // device scalars of this solve (k_dim + 2 doubles behind the work vectors)
   double *h_sc;      // host mirror (pinned)

   int fetch(int cnt)
   {
      CUDA_TRY(cudaMemcpyAsync(h_sc, d_sc, sizeof(double) * (size_t)cnt, cudaMemcpyDeviceToHost, c->stream));
      if (c->use_p2p) { CUDA_TRY(cudaMemcpyAsync(c->h_err, c->d_err, sizeof(int), cudaMemcpyDeviceToHost, c->stream)); }
      CUDA_TRY(cudaStreamSynchronize(c->stream));
      if (c->use_p2p && *c->h_err)
      {
         b200_set_error("peer-memory exchange timed out (a neighbour rank did not publish its halo)");
         return B200_ERR_NCCL;
      }
      return B200_OK;
   }
   int matvec(double alpha, const double *x, double beta, const double *b, double *y)
   {
      return op_matvec(A, alpha, x, beta, b, y);
   }
   int matvecT(double alpha, const double *x, double beta, double *y) { return op_matvecT(A, alpha, x, beta, y); }
   int precond(const double *r, double *z, const double **zr)
   {
      *zr = z;
      if (amg)
      {
         double *out = nullptr;
         B200_TRY(amg_precond_ptr(amg, r, z, &out));
         *zr = out;
         return B200_OK;
      }
      if (diag_scale)
      {
         FDiagScale f{A->d_diagonal, r, z};
         return run_ewise(c, n, f, nullptr);
      }
      *zr = r;       // hypre_ParKrylovIdentity
      return B200_OK;
   }
   int copy(const double *x, double *y) { return b200_vec_copy(c, n, x, y); }
   int scale(double a, double *y) { return b200_vec_scale(c, n, a, y); }
   int axpy(double a, const double *x, double *y) { return vec_axpy(c, n, a, x, y); }
   int dots(int k, const double *const *xs, const double *const *ys, double *out)
   {
      for (int q = 0; q < k; q++) { B200_TRY(vec_dot(c, n, xs[q], ys[q], d_sc + q)); }
      B200_TRY(scalars_allreduce(c, d_sc, k));
      B200_TRY(fetch(k));
      for (int q = 0; q < k; q++) { out[q] = h_sc[q]; }
      return B200_OK;
   }
   int mgs(int i, double *const *p, double *h)
   {
      B200_TRY(vec_dot(c, n, p[0], p[i], d_sc + 0));
      B200_TRY(scalars_allreduce(c, d_sc + 0, 1));
      for (int j = 0; j < i; j++)
      {
         FMgsStep f{d_sc + j, p[j], j + 1 < i ? p[j + 1] : p[i], p[i], 0.0};
         B200_TRY(run_ewise(c, n, f, d_sc + j + 1));
         B200_TRY(scalars_allreduce(c, d_sc + j + 1, 1));
      }
      B200_TRY(fetch(i + 1));
      for (int j = 0; j <= i; j++) { h[j] = h_sc[j]; }
      return B200_OK;
   }
   // COGMRES: ceil(i/4) passes for the i inner products (one all-reduce for all of them), ceil(i/4) passes for
   // the update, the last of which also yields <p_i, p_i>; d_sc[0..i-1] = h, d_sc[i] = the squared norm
   int cgs(int i, double *const *p, double *h)
   {
      for (int j0 = 0; j0 < i; j0 += 4)
      {
         const int cnt = i - j0 < 4 ? i - j0 : 4;
         FMassDot4 f{cnt, p[i], {p[j0], p[j0 + (cnt > 1 ? 1 : 0)], p[j0 + (cnt > 2 ? 2 : 0)], p[j0 + (cnt > 3 ? 3 : 0)]}};
         // the reduction writes NRED = 4 scalars: d_sc has k_dim + 2 + 4 entries, later passes overwrite the spill
         B200_TRY(run_ewise(c, n, f, d_sc + j0));
      }
      for (int j0 = 0; j0 < i; j0 += AR_CHUNK) { B200_TRY(scalars_allreduce(c, d_sc + j0, i - j0 < AR_CHUNK ? i - j0 : AR_CHUNK)); }
      for (int j0 = 0; j0 < i; j0 += 4)
      {
         const int cnt = i - j0 < 4 ? i - j0 : 4;
         const int last = j0 + 4 >= i;
         FMassAxpy4 f{cnt, last, d_sc + j0, {p[j0], p[j0 + (cnt > 1 ? 1 : 0)], p[j0 + (cnt > 2 ? 2 : 0)], p[j0 + (cnt > 3 ? 3 : 0)]},
                      p[i], {0.0, 0.0, 0.0, 0.0}};
         // only the last pass reduces; the others park their (zero) sum in the spare slot behind the column
         B200_TRY(run_ewise(c, n, f, d_sc + (last ? i : i + 1)));
      }
      B200_TRY(scalars_allreduce(c, d_sc + i, 1));
      B200_TRY(fetch(i + 1));
      for (int j = 0; j <= i; j++) { h[j] = h_sc[j]; }
      return B200_OK;
   }
   int bicg_half(double a, const double *v, const double *q, double *x, double *r)
   {
      FBicgHalf f{a, v, q, x, r};
      return run_ewise(c, n, f, nullptr);
   }
   int bicg_full(double g, const double *v, const double *s, const double *r0, double *x, double *r, double *out)
   {
      FBicgFull f{g, v, s, r0, x, r};
      B200_TRY(run_ewise(c, n, f, d_sc));
      B200_TRY(scalars_allreduce(c, d_sc, 2));
      B200_TRY(fetch(2));
      out[0] = h_sc[0]; out[1] = h_sc[1];
      return B200_OK;
   }
   int bicg_dir(double g, double f_, const double *q, const double *r, double *p)
   {
      FBicgDir f{g, f_, q, r, p};
      return run_ewise(c, n, f, nullptr);
   }
};

// work vectors of one solve: `nvec` vectors of n doubles (each starting on a 256-byte boundary) followed by
// `nsc` device scalars, all in one cached allocation of the context
static int krylov_workspace(b200_ctx *c, int n, int nvec, int nsc, std::vector<double *> *vec, double **d_sc,
                            double **h_sc)
{
   const size_t stride = ((size_t)(n > 0 ? n : 1) + 31) / 32 * 32;
   double *base = nullptr;
   B200_TRY(ctx_workspace(c, WS_KRYLOV, stride * (size_t)nvec + (size_t)nsc, &base));
   vec->resize(nvec);
   for (int k = 0; k < nvec; k++) { (*vec)[k] = base + stride * (size_t)k; }
   *d_sc = base + stride * (size_t)nvec;
   if (c->h_kry_n < (size_t)nsc)
   {
      if (c->h_kry) { CUDA_TRY(cudaFreeHost(c->h_kry)); c->h_kry = nullptr; c->h_kry_n = 0; }
      CUDA_TRY(cudaMallocHost((void **)&c->h_kry, sizeof(double) * (size_t)nsc));
      c->h_kry_n = (size_t)nsc;
   }
   *h_sc = c->h_kry;
   return B200_OK;
}

static int krylov_prepare(b200_ctx *c, b200_parcsr *A, b200_amg *amg, int diag_scale)
{
   if (!amg && diag_scale) { B200_TRY(parcsr_ensure_L(A)); }
   B200_TRY(ctx_rendezvous(c));
   return B200_OK;
}

// drain the stream and surface an exchange time-out that no later fetch would see
static int krylov_finish(b200_ctx *c, int status)
{
   cudaError_t e = cudaStreamSynchronize(c->stream);
   if (status == B200_OK && e != cudaSuccess) { CUDA_TRY(e); }
   if (status == B200_OK && c->use_p2p && c->h_err && *c->h_err) { status = B200_ERR_NCCL; }
   return status;
}

extern "C" int b200_gmres_solve(b200_ctx *c, b200_parcsr *A, b200_amg *amg, const b200_gmres_params *prm,
                                const double *b, double *x, b200_krylov_result *res, double *norms_h)
{
   ARG_CHECK(c != nullptr && A != nullptr && prm != nullptr && b != nullptr && x != nullptr && res != nullptr);
   ARG_CHECK(prm->max_iter >= 0 && prm->k_dim >= 1);
   const int n = A->n_rows;
   std::vector<double *> vec;
   double *d_sc = nullptr, *h_sc = nullptr;
   ARG_CHECK(prm->ortho == 0 || prm->ortho == 1);
   // flexible GMRES keeps k_dim preconditioned vectors behind p[0..k_dim], r, w
   B200_TRY(krylov_workspace(c, n, prm->k_dim + 3 + (prm->flex ? prm->k_dim : 0), prm->k_dim + 2 + 4, &vec, &d_sc, &h_sc));
   B200_TRY(krylov_prepare(c, A, amg, prm->diag_scale));
   DevKrylovOps ops{c, A, amg, prm->diag_scale, n, d_sc, h_sc};
   // vec[0..k_dim] = the Arnoldi basis p, then r and w (gmres.c:203-240)
   const int status = gmres_core(ops, prm, b, x, vec.data(), vec[prm->k_dim + 1], vec[prm->k_dim + 2], res, norms_h,
                                 prm->flex ? vec.data() + prm->k_dim + 3 : nullptr);
   return krylov_finish(c, status);
}

extern "C" int b200_bicgstab_solve(b200_ctx *c, b200_parcsr *A, b200_amg *amg, const b200_bicgstab_params *prm,
                                   const double *b, double *x, b200_krylov_result *res, double *norms_h)
{
   ARG_CHECK(c != nullptr && A != nullptr && prm != nullptr && b != nullptr && x != nullptr && res != nullptr);
   ARG_CHECK(prm->max_iter >= 0);
   const int n = A->n_rows;
   std::vector<double *> vec;
   double *d_sc = nullptr, *h_sc = nullptr;
   B200_TRY(krylov_workspace(c, n, 6, 4, &vec, &d_sc, &h_sc));
   B200_TRY(krylov_prepare(c, A, amg, prm->diag_scale));
   DevKrylovOps ops{c, A, amg, prm->diag_scale, n, d_sc, h_sc};
   // r, r0, s, v, p, q (bicgstab.c:170-193)
   const int status = bicgstab_core(ops, prm, b, x, vec[0], vec[1], vec[2], vec[3], vec[4], vec[5], res, norms_h);
   return krylov_finish(c, status);
}

extern "C" int b200_lgmres_solve(b200_ctx *c, b200_parcsr *A, b200_amg *amg, const b200_lgmres_params *prm,
                                 const double *b, double *x, b200_krylov_result *res, double *norms_h)
{
   ARG_CHECK(c != nullptr && A != nullptr && prm != nullptr && b != nullptr && x != nullptr && res != nullptr);
   ARG_CHECK(prm->max_iter >= 0 && prm->aug_dim >= 0 && prm->k_dim > prm->aug_dim);
   const int n = A->n_rows;
   const int kd = prm->k_dim, ad = prm->aug_dim;
   std::vector<double *> vec;
   double *d_sc = nullptr, *h_sc = nullptr;
   // p[0..k_dim], aug_vecs[0..aug_dim], a_aug_vecs[0..aug_dim-1], r, w (lgmres.c:255-290)
   const int nvec = (kd + 1) + (ad + 1) + ad + 2;
   B200_TRY(krylov_workspace(c, n, nvec, kd + ad + 2 + 4, &vec, &d_sc, &h_sc));
   B200_TRY(krylov_prepare(c, A, amg, prm->diag_scale));
   DevKrylovOps ops{c, A, amg, prm->diag_scale, n, d_sc, h_sc};
   const int status = lgmres_core(ops, prm, b, x, vec.data(), vec.data() + kd + 1, vec.data() + kd + 1 + ad + 1,
                                  vec[nvec - 2], vec[nvec - 1], res, norms_h);
   return krylov_finish(c, status);
}

// host b/x: same staging as b200_pcg_solve_host
template <class Prm, class Solve>
static int solve_host(b200_ctx *c, b200_parcsr *A, b200_amg *amg, const Prm *prm, const double *b_h, double *x_h,
                      b200_krylov_result *res, double *norms_h, Solve solve)
{
   ARG_CHECK(c != nullptr && A != nullptr && b_h != nullptr && x_h != nullptr);
   const int n = A->n_rows;
   double *b = nullptr, *x = nullptr;
   B200_TRY(ctx_workspace(c, WS_HOST_B, n, &b));
   B200_TRY(ctx_workspace(c, WS_HOST_X, n, &x));
   CUDA_TRY(cudaMemcpyAsync(b, b_h, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
   CUDA_TRY(cudaMemcpyAsync(x, x_h, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
   B200_TRY(solve(c, A, amg, prm, b, x, res, norms_h));
   return b200_d2h(c, x_h, x, n);
}

extern "C" int b200_gmres_solve_host(b200_ctx *c, b200_parcsr *A, b200_amg *amg, const b200_gmres_params *prm,
                                     const double *b_h, double *x_h, b200_krylov_result *res, double *norms_h)
{
   return solve_host(c, A, amg, prm, b_h, x_h, res, norms_h, b200_gmres_solve);
}

extern "C" int b200_bicgstab_solve_host(b200_ctx *c, b200_parcsr *A, b200_amg *amg, const b200_bicgstab_params *prm,
                                        const double *b_h, double *x_h, b200_krylov_result *res, double *norms_h)
{
   return solve_host(c, A, amg, prm, b_h, x_h, res, norms_h, b200_bicgstab_solve);
}

extern "C" int b200_lgmres_solve_host(b200_ctx *c, b200_parcsr *A, b200_amg *amg, const b200_lgmres_params *prm,
                                      const double *b_h, double *x_h, b200_krylov_result *res, double *norms_h)
{
   return solve_host(c, A, amg, prm, b_h, x_h, res, norms_h, b200_lgmres_solve);
}

