// krylov_host_backend.cpp -- TEST INFRASTRUCTURE (never linked into libb200amg.so or libHYPRE_b200.so).
//
// Instantiates the GMRES / BiCGSTAB loops of hypre_b200/csrc/krylov_core.hpp -- the same templates the device
// library instantiates over its CUDA backend -- over a plain host backend whose operator and preconditioner
// are the oracle's (orc_csr_matvec, orc_amg_cycle).  tests/test_krylov_core_host.py compares the iterates with
// orc_gmres / orc_bicgstab, which are pinned to the compiled reference: the scalar logic of the loops
// (Givens rotations, restarts, true-residual checks, breakdown exits) is thereby checked without a GPU.  The
// fused methods below are also the specification of the device functors of krylov.cu.
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../hypre_b200/csrc/krylov_core.hpp"
#include "../../oracle/amg_oracle.h"

struct HostKrylovOps {
   int n; const int *ia; const int *ja; const double *a; orc_amg *amg; int kind;   // kind: 0 AMG, 1 diag, 2 identity
   long calls = 0;
   int matvec(double alpha, const double *x, double beta, const double *b, double *y)
   {
      calls++;
      orc_csr_matvec(n, ia, ja, a, alpha, x, beta, b, y);
      return B200_OK;
   }
   int matvecT(double alpha, const double *x, double beta, double *y)
   {
      orc_csr_matvecT(n, n, ia, ja, a, alpha, x, beta, y);
      return B200_OK;
   }
   int precond(const double *r, double *z, const double **zr)
   {
      *zr = z;
      if (kind == 0) { memset(z, 0, sizeof(double) * (size_t)n); orc_amg_cycle(amg, r, z, 1); }
      else if (kind == 1) { for (int k = 0; k < n; k++) { z[k] = r[k] / a[ia[k]]; } }
      else { *zr = r; }
      return B200_OK;
   }
   int copy(const double *x, double *y) { if (x != y) { memcpy(y, x, sizeof(double) * (size_t)n); } return B200_OK; }
   int scale(double s, double *y)
   {
      if (s == 1.0) { return B200_OK; }
      for (int k = 0; k < n; k++) { y[k] *= s; }
      return B200_OK;
   }
   int axpy(double s, const double *x, double *y) { for (int k = 0; k < n; k++) { y[k] += s * x[k]; } return B200_OK; }
   static double dot(int n, const double *x, const double *y)
   {
      double s = 0.0;
      for (int k = 0; k < n; k++) { s += x[k] * y[k]; }
      return s;
   }
   int dots(int k, const double *const *xs, const double *const *ys, double *out)
   {
      for (int q = 0; q < k; q++) { out[q] = dot(n, xs[q], ys[q]); }
      return B200_OK;
   }
   int mgs(int i, double *const *p, double *h)
   {
      for (int j = 0; j < i; j++)
      {
         h[j] = dot(n, p[j], p[i]);
         for (int k = 0; k < n; k++) { p[i][k] += -h[j] * p[j][k]; }
      }
      h[i] = dot(n, p[i], p[i]);
      return B200_OK;
   }
   int cgs(int i, double *const *p, double *h)
   {
      for (int j = 0; j < i; j++) { h[j] = dot(n, p[i], p[j]); }
      for (int j = 0; j < i; j++) { for (int k = 0; k < n; k++) { p[i][k] += -h[j] * p[j][k]; } }
      h[i] = dot(n, p[i], p[i]);
      return B200_OK;
   }
   int bicg_half(double al, const double *v, const double *q, double *x, double *r)
   {
      for (int k = 0; k < n; k++) { x[k] += al * v[k]; }
      for (int k = 0; k < n; k++) { r[k] += -al * q[k]; }
      return B200_OK;
   }
   int bicg_full(double g, const double *v, const double *s, const double *r0, double *x, double *r, double *out)
   {
      for (int k = 0; k < n; k++) { x[k] += g * v[k]; }
      for (int k = 0; k < n; k++) { r[k] += -g * s[k]; }
      out[0] = dot(n, r, r); out[1] = dot(n, r0, r);
      return B200_OK;
   }
   int bicg_dir(double g, double f, const double *q, const double *r, double *p)
   {
      for (int k = 0; k < n; k++) { p[k] += -g * q[k]; }
      for (int k = 0; k < n; k++) { p[k] *= f; }
      for (int k = 0; k < n; k++) { p[k] += r[k]; }
      return B200_OK;
   }
};

extern "C" int kht_gmres(int n, const int *ia, const int *ja, const double *a, orc_amg *amg, int kind,
                         const b200_gmres_params *prm, const double *b, double *x, b200_krylov_result *res,
                         double *norms)
{
   HostKrylovOps ops{n, ia, ja, a, amg, kind};
   std::vector<std::vector<double>> store(prm->k_dim + 3 + (prm->flex ? prm->k_dim : 0), std::vector<double>((size_t)n, 0.0));
   std::vector<double *> vec;
   for (auto &v : store) { vec.push_back(v.data()); }
   return gmres_core(ops, prm, b, x, vec.data(), vec[prm->k_dim + 1], vec[prm->k_dim + 2], res, norms,
                     prm->flex ? vec.data() + prm->k_dim + 3 : nullptr);
}

extern "C" int kht_bicgstab(int n, const int *ia, const int *ja, const double *a, orc_amg *amg, int kind,
                            const b200_bicgstab_params *prm, const double *b, double *x, b200_krylov_result *res,
                            double *norms)
{
   HostKrylovOps ops{n, ia, ja, a, amg, kind};
   std::vector<std::vector<double>> store(6, std::vector<double>((size_t)n, 0.0));
   return bicgstab_core(ops, prm, b, x, store[0].data(), store[1].data(), store[2].data(), store[3].data(),
                        store[4].data(), store[5].data(), res, norms);
}

// ---------------------------------------------------------------- N-rank form of the host backend
// The same loops on a row-partitioned problem: the operator (halo exchange + local rows) and the sum of the
// inner products over the ranks are callbacks supplied by the test (tests/dist_krylov_worker.py: gloo), which
// is how the device backend is organised as well (op_matvec does the exchange, scalars_allreduce the sums).
typedef void (*kht_matvec_cb)(double alpha, const double *x, double beta, const double *b, double *y);
typedef void (*kht_allreduce_cb)(double *vals, int cnt);

struct CbKrylovOps {
   int n; const double *diag; kht_matvec_cb mv; kht_allreduce_cb ar;     // diag == nullptr: no preconditioner
   int matvec(double alpha, const double *x, double beta, const double *b, double *y) { mv(alpha, x, beta, b, y); return B200_OK; }
   int precond(const double *r, double *z, const double **zr)
   {
      *zr = z;
      if (diag) { for (int k = 0; k < n; k++) { z[k] = r[k] / diag[k]; } }
      else { *zr = r; }
      return B200_OK;
   }
   int copy(const double *x, double *y) { if (x != y) { memcpy(y, x, sizeof(double) * (size_t)n); } return B200_OK; }
   int scale(double s, double *y) { if (s != 1.0) { for (int k = 0; k < n; k++) { y[k] *= s; } } return B200_OK; }
   int axpy(double s, const double *x, double *y) { for (int k = 0; k < n; k++) { y[k] += s * x[k]; } return B200_OK; }
   double ldot(const double *x, const double *y) const { return HostKrylovOps::dot(n, x, y); }
   int dots(int k, const double *const *xs, const double *const *ys, double *out)
   {
      for (int q = 0; q < k; q++) { out[q] = ldot(xs[q], ys[q]); }
      ar(out, k);
      return B200_OK;
   }
   int mgs(int i, double *const *p, double *h)
   {
      for (int j = 0; j < i; j++)
      {
         h[j] = ldot(p[j], p[i]); ar(h + j, 1);
         for (int k = 0; k < n; k++) { p[i][k] += -h[j] * p[j][k]; }
      }
      h[i] = ldot(p[i], p[i]); ar(h + i, 1);
      return B200_OK;
   }
   int cgs(int i, double *const *p, double *h)
   {
      for (int j = 0; j < i; j++) { h[j] = ldot(p[i], p[j]); }
      if (i > 0) { ar(h, i); }
      for (int j = 0; j < i; j++) { for (int k = 0; k < n; k++) { p[i][k] += -h[j] * p[j][k]; } }
      h[i] = ldot(p[i], p[i]); ar(h + i, 1);
      return B200_OK;
   }
   int bicg_half(double al, const double *v, const double *q, double *x, double *r)
   {
      for (int k = 0; k < n; k++) { x[k] += al * v[k]; }
      for (int k = 0; k < n; k++) { r[k] += -al * q[k]; }
      return B200_OK;
   }
   int bicg_full(double g, const double *v, const double *s, const double *r0, double *x, double *r, double *out)
   {
      for (int k = 0; k < n; k++) { x[k] += g * v[k]; }
      for (int k = 0; k < n; k++) { r[k] += -g * s[k]; }
      out[0] = ldot(r, r); out[1] = ldot(r0, r);
      ar(out, 2);
      return B200_OK;
   }
   int bicg_dir(double g, double f, const double *q, const double *r, double *p)
   {
      for (int k = 0; k < n; k++) { p[k] += -g * q[k]; }
      for (int k = 0; k < n; k++) { p[k] *= f; }
      for (int k = 0; k < n; k++) { p[k] += r[k]; }
      return B200_OK;
   }
};

extern "C" int kht_gmres_dist(int n_local, const double *diag, kht_matvec_cb mv, kht_allreduce_cb ar,
                              const b200_gmres_params *prm, const double *b, double *x, b200_krylov_result *res,
                              double *norms)
{
   CbKrylovOps ops{n_local, diag, mv, ar};
   std::vector<std::vector<double>> store(prm->k_dim + 3 + (prm->flex ? prm->k_dim : 0),
                                          std::vector<double>((size_t)(n_local > 0 ? n_local : 1), 0.0));
   std::vector<double *> vec;
   for (auto &v : store) { vec.push_back(v.data()); }
   return gmres_core(ops, prm, b, x, vec.data(), vec[prm->k_dim + 1], vec[prm->k_dim + 2], res, norms,
                     prm->flex ? vec.data() + prm->k_dim + 3 : nullptr);
}

extern "C" int kht_bicgstab_dist(int n_local, const double *diag, kht_matvec_cb mv, kht_allreduce_cb ar,
                                 const b200_bicgstab_params *prm, const double *b, double *x, b200_krylov_result *res,
                                 double *norms)
{
   CbKrylovOps ops{n_local, diag, mv, ar};
   std::vector<std::vector<double>> store(6, std::vector<double>((size_t)(n_local > 0 ? n_local : 1), 0.0));
   return bicgstab_core(ops, prm, b, x, store[0].data(), store[1].data(), store[2].data(), store[3].data(),
                        store[4].data(), store[5].data(), res, norms);
}

extern "C" int kht_lgmres(int n, const int *ia, const int *ja, const double *a, orc_amg *amg, int kind,
                          const b200_lgmres_params *prm, const double *b, double *x, b200_krylov_result *res,
                          double *norms)
{
   HostKrylovOps ops{n, ia, ja, a, amg, kind};
   const int nv = prm->k_dim + 1 + (prm->aug_dim + 1) + prm->aug_dim + 2;
   std::vector<std::vector<double>> store(nv, std::vector<double>((size_t)n, 0.0));
   std::vector<double *> vec;
   for (auto &v : store) { vec.push_back(v.data()); }
   double *const *p = vec.data();
   double *const *av = vec.data() + prm->k_dim + 1;
   double *const *aav = av + prm->aug_dim + 1;
   return lgmres_core(ops, prm, b, x, p, av, aav, vec[nv - 2], vec[nv - 1], res, norms);
}

extern "C" int kht_lgmres_dist(int n_local, const double *diag, kht_matvec_cb mv, kht_allreduce_cb ar,
                               const b200_lgmres_params *prm, const double *b, double *x, b200_krylov_result *res,
                               double *norms)
{
   CbKrylovOps ops{n_local, diag, mv, ar};
   const int nv = prm->k_dim + 1 + (prm->aug_dim + 1) + prm->aug_dim + 2;
   std::vector<std::vector<double>> store(nv, std::vector<double>((size_t)(n_local > 0 ? n_local : 1), 0.0));
   std::vector<double *> vec;
   for (auto &v : store) { vec.push_back(v.data()); }
   double *const *av = vec.data() + prm->k_dim + 1;
   return lgmres_core(ops, prm, b, x, vec.data(), av, av + prm->aug_dim + 1, vec[nv - 2], vec[nv - 1], res, norms);
}

