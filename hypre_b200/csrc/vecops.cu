// vecops.cu -- BLAS-1 on device (hypre_SeqVector* of src/seq_mv/vector.c).
#include "vecops.cuh"
#include "p2p.cuh"

struct FFill  { static constexpr int NRED = 0; __device__ void init() {} double a; double *y;
   __device__ void operator()(int i, double *) const { y[i] = a; } };
struct FScale { static constexpr int NRED = 0; __device__ void init() {} double a; double *y;
   __device__ void operator()(int i, double *) const { y[i] *= a; } };
struct FAxpy  { static constexpr int NRED = 0; __device__ void init() {} double a; const double *x; double *y;
   __device__ void operator()(int i, double *) const { y[i] = fma(a, x[i], y[i]); } };
struct FDot   { static constexpr int NRED = 1; __device__ void init() {} const double *x; const double *y;
   __device__ void operator()(int i, double *red) const { red[0] = fma(x[i], y[i], red[0]); } };

static int tail_blas1(b200_ctx *c, int kind, int n, double a, const double *x, double *y)
{
   TailOp op;
   memset(&op, 0, sizeof(op));
   op.kind = kind; op.n = n; op.a = a; op.x = x; op.y = y;
   return tail_append(c, op);
}

int vec_fill(b200_ctx *c, int n, double a, double *y)
{
   if (n == 0) { return B200_OK; }
   if (tail_recording(c)) { return tail_blas1(c, TK_FILL, n, a, nullptr, y); }
   if (a == 0.0) { CUDA_TRY(cudaMemsetAsync(y, 0, sizeof(double) * (size_t)n, c->stream)); return B200_OK; }
   FFill f{a, y};
   return run_ewise(c, n, f, nullptr);
}
int vec_scale(b200_ctx *c, int n, double a, double *y)
{
   if (tail_recording(c)) { return n > 0 ? tail_blas1(c, TK_SCALE, n, a, nullptr, y) : B200_OK; }
   FScale f{a, y};
   return run_ewise(c, n, f, nullptr);
}
int vec_axpy(b200_ctx *c, int n, double a, const double *x, double *y)
{
   if (tail_recording(c)) { return n > 0 ? tail_blas1(c, TK_AXPY, n, a, x, y) : B200_OK; }
   FAxpy f{a, x, y};
   return run_ewise(c, n, f, nullptr);
}
int vec_dot(b200_ctx *c, int n, const double *x, const double *y, double *d_out)
{
   FDot f{x, y};
   return run_ewise(c, n, f, d_out);
}

int scalars_allreduce(b200_ctx *c, double *d, int cnt)
{
   if (c->nranks <= 1 || cnt == 0) { return B200_OK; }
   if (c->use_p2p) { return p2p_allreduce(c, d, cnt); }
   // hypre_MPI_Allreduce of hypre_ParVectorInnerProd (src/parcsr_mv/par_vector.c:526)
   NCCL_TRY(ncclAllReduce(d, d, (size_t)cnt, ncclDouble, ncclSum, c->comm, c->stream));
   return B200_OK;
}

int scalars_to_host(b200_ctx *c, int first, int cnt)
{
   CUDA_TRY(cudaMemcpyAsync(c->h_scalars + first, c->d_scalars + first, sizeof(double) * (size_t)cnt,
                            cudaMemcpyDeviceToHost, c->stream));
   if (c->use_p2p) { CUDA_TRY(cudaMemcpyAsync(c->h_err, c->d_err, sizeof(int), cudaMemcpyDeviceToHost, c->stream)); }
   CUDA_TRY(cudaStreamSynchronize(c->stream));
   if (c->use_p2p && *c->h_err)
   {
      b200_set_error("peer-memory exchange timed out (a neighbour rank did not publish its halo)");
      return B200_ERR_NCCL;
   }
   return B200_OK;
}
