// vecops.cuh -- streaming elementwise kernels with fused reductions.
#pragma once
#include "spmv.cuh"
#include "rowlaunch.cuh"

// f(i, red) is applied to every i in [0, n); 4 independent elements per thread
// per trip keep enough bytes in flight for HBM3e.
template <class F>
__global__ void __launch_bounds__(ROWK_BLOCK) ewise_kernel(int n, F f, RedCtx rc)
{
   RETURN_IF_STOPPED(rc);
   f.init();      // per-thread preparation (scalars read once from device memory); no-op for most functors
   double red[F::NRED > 0 ? F::NRED : 1];
#pragma unroll
   for (int k = 0; k < (F::NRED > 0 ? F::NRED : 1); k++) { red[k] = 0.0; }
   const int stride = gridDim.x * blockDim.x;
   int i = blockIdx.x * blockDim.x + threadIdx.x;
   for (; i + 3 * stride < n; i += 4 * stride)
   {
      f(i, red); f(i + stride, red); f(i + 2 * stride, red); f(i + 3 * stride, red);
   }
   for (; i < n; i += stride) { f(i, red); }
   grid_reduce<F::NRED>(red, rc);
}

template <class F>
static int run_ewise(b200_ctx *ctx, int n, const F &f, double *red_out)
{
   if (tail_recording(ctx))
   {
      if (n == 0) { return B200_OK; }
      TailOp op;
      memset(&op, 0, sizeof(op));
      op.n = n;
      return tail_record_fn(ctx, op, f);
   }
   RedCtx rc = make_redctx(ctx, red_out, 0);
   if (n == 0)
   {
      if (F::NRED > 0) { CUDA_TRY(cudaMemsetAsync(red_out, 0, sizeof(double) * F::NRED, ctx->stream)); }
      return B200_OK;
   }
   int grid = rows_grid(((long long)n + 3) / 4, ROWK_BLOCK);
   LAUNCH(ctx, (ewise_kernel<F>), grid, ROWK_BLOCK, 0, n, f, rc);
   CHECK_LAUNCH();
   return B200_OK;
}

int vec_fill(b200_ctx *c, int n, double a, double *y);
int vec_scale(b200_ctx *c, int n, double a, double *y);
int vec_axpy(b200_ctx *c, int n, double a, const double *x, double *y);
int vec_dot(b200_ctx *c, int n, const double *x, const double *y, double *d_out);
// sum the device scalars d[0..cnt) over all ranks (no-op for one rank)
int scalars_allreduce(b200_ctx *c, double *d, int cnt);
// copy d_scalars[first..first+cnt) into the pinned mirror and wait for it
int scalars_to_host(b200_ctx *c, int first, int cnt);
