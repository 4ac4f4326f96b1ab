// ops.cuh -- internal operator entry points shared by amg.cu / pcg.cu / the C ABI.
#pragma once
#include "common.cuh"

int parcsr_ensure_T(b200_parcsr *A);
int parcsr_ensure_L(b200_parcsr *A);
int halo_reverse_start(b200_parcsr *A);
int halo_reverse_finish_add(b200_parcsr *A, double alpha, double *y);

// A dot product fused into the last launch of an operator: out[0] = <with, result> (local part), *done is set
// when the operator took the request (otherwise the caller reduces in a pass of its own)
struct FusedDot { const double *with; double *out; bool *done; };

// y = alpha*A*x + beta*b
int op_matvec(b200_parcsr *A, double alpha, const double *x, double beta, const double *b, double *y);
// y = A*x and d_out[0] = <y, x> (local part) in the same launch  [PCG: s = A p, <s,p>]
int op_matvec_dot(b200_parcsr *A, const double *x, double *y, double *d_out);
// gathered operands can be composite (formed on the fly by the epilogue's xval) on this operator
bool op_can_fuse_gather(const b200_parcsr *A);
// PCG: p_new = z + beta p_old (buffers chosen by the parity of *d_iter), s = A p_new, d_out[0] = <s, p_new>
int op_matvec_dot_direction(b200_parcsr *A, const double *d_beta, const int *d_iter, const double *z,
                            double *pbuf0, double *pbuf1, double *s, double *d_out);
// y = alpha*A^T*x + beta*y
int op_matvecT(b200_parcsr *A, double alpha, const double *x, double beta, double *y);
// u_out = u_in + w*(f - A u_in)./l1          (zero_guess: u_out = w*f./l1, u_in unused)
int op_l1jacobi(b200_parcsr *A, const double *f, const double *l1, double w,
                const double *u_in, double *u_out, int zero_guess, const FusedDot *fd = nullptr,
                const int *marker = nullptr, int points = 0);   // marker / points: CF-ordered sweep, rows with marker != points keep u
// Chebyshev, in place on u; v, r, tmp are work vectors
int op_cheby(b200_parcsr *A, const double *f, const double *ds, const double *coefs_h, int order,
             int scale, double *u, double *v, double *r, double *tmp, int zero_guess, const FusedDot *fd = nullptr);
// two-stage Gauss-Seidel, in place on u; z, tmp work vectors
int op_gs2(b200_parcsr *A, const double *f, double w, int inner, double *u, double *z, double *tmp,
           int zero_guess, const FusedDot *fd = nullptr);
// dense coarse solve u = Ainv * f_all (one block per 8 rows)
int op_dense_matvec(b200_ctx *ctx, int n_rows, int n_cols, const double *Ainv_rows, const double *f_all,
                    double *u);
