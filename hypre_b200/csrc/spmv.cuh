// spmv.cuh -- the fused row kernels.  Every solve-phase operator that touches a
// sparse matrix is "row sum t_i = sum_j a_ij x_j, then an epilogue on (i, t_i)":
// SpMV, residual, l1-Jacobi, the Chebyshev steps, two-stage GS, restriction
// (with the stored transpose) and prolongation.  The matrix is streamed once,
// the vectors the epilogue needs are streamed once, and up to B200_MAX_RED dot
// products are reduced in the same launch.
#pragma once
#include "common.cuh"

constexpr int ROWK_BLOCK = 256;

// streaming (read-once) loads: keep L1 for the gathered x
__device__ __forceinline__ double ld_stream(const double *p)
{
   double v;
   asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
   return v;
}
__device__ __forceinline__ int ld_stream(const int *p)
{
   int v;
   asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(v) : "l"(p));
   return v;
}

// ---------------------------------------------------------------- reductions
struct RedCtx {
   double   *partials;    // [B200_MAX_RED][B200_MAX_PARTIALS]
   unsigned *counter;
   double   *out;         // device scalars, NRED consecutive
   int       accumulate;  // add to out[] instead of overwriting (second pass of a split launch)
};

// Deterministic grid reduction: block partials in a fixed slot, last block to
// finish sums them in index order.  gridDim.x <= B200_MAX_PARTIALS is
// guaranteed by the launchers (they fold the grid with a stride loop).
template <int NRED>
__device__ __forceinline__ void grid_reduce(double (&red)[NRED > 0 ? NRED : 1], const RedCtx &rc)
{
   if (NRED == 0) { return; }
   __shared__ double s_red[(NRED > 0 ? NRED : 1)][ROWK_BLOCK / 32];
   __shared__ bool   s_last;
   const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
   for (int k = 0; k < NRED; k++)
   {
      double v = red[k];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) { v += __shfl_down_sync(0xffffffffu, v, o); }
      if (lane == 0) { s_red[k][warp] = v; }
   }
   __syncthreads();
   if (threadIdx.x == 0)
   {
#pragma unroll
      for (int k = 0; k < NRED; k++)
      {
         double v = 0.0;
         for (int w = 0; w < (int)(blockDim.x >> 5); w++) { v += s_red[k][w]; }
         rc.partials[(size_t)k * B200_MAX_PARTIALS + blockIdx.x] = v;
      }
      __threadfence();
      unsigned done = atomicAdd(rc.counter, 1u);
      s_last = (done == gridDim.x - 1);
   }
   __syncthreads();
   if (s_last)
   {
      __threadfence();
#pragma unroll
      for (int k = 0; k < NRED; k++)
      {
         // fixed-order tree over the block partials
         double v = 0.0;
         for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x)
         {
            v += ((volatile double *)rc.partials)[(size_t)k * B200_MAX_PARTIALS + b];
         }
#pragma unroll
         for (int o = 16; o > 0; o >>= 1) { v += __shfl_down_sync(0xffffffffu, v, o); }
         __syncthreads();
         if (lane == 0) { s_red[k][warp] = v; }
         __syncthreads();
         if (threadIdx.x == 0)
         {
            double t = 0.0;
            for (int w = 0; w < (int)(blockDim.x >> 5); w++) { t += s_red[k][w]; }
            rc.out[k] = rc.accumulate ? rc.out[k] + t : t;
         }
      }
      if (threadIdx.x == 0) { *rc.counter = 0u; }
   }
}

// ---------------------------------------------------------------- row sums
// sliced-ELL, one thread per row (lane = row within the 32-row slice)
__device__ __forceinline__ double rowsum_sell(const MatView &D, int slice, int lane,
                                              const double *__restrict__ x)
{
   const long long off = D.sl_off[slice];
   const int width = (int)((D.sl_off[slice + 1] - off) >> 5);
   const int    *ci = D.sl_ci + off + lane;
   const double *va = D.sl_va + off + lane;
   double t = 0.0;
   int k = 0;
   for (; k + 4 <= width; k += 4)
   {
      int    c0 = ld_stream(ci + (k + 0) * 32), c1 = ld_stream(ci + (k + 1) * 32);
      int    c2 = ld_stream(ci + (k + 2) * 32), c3 = ld_stream(ci + (k + 3) * 32);
      double v0 = ld_stream(va + (k + 0) * 32), v1 = ld_stream(va + (k + 1) * 32);
      double v2 = ld_stream(va + (k + 2) * 32), v3 = ld_stream(va + (k + 3) * 32);
      double x0 = __ldg(x + c0), x1 = __ldg(x + c1), x2 = __ldg(x + c2), x3 = __ldg(x + c3);
      t = fma(v0, x0, t); t = fma(v1, x1, t); t = fma(v2, x2, t); t = fma(v3, x3, t);
   }
   for (; k < width; k++)
   {
      int    c = ld_stream(ci + k * 32);
      double v = ld_stream(va + k * 32);
      t = fma(v, __ldg(x + c), t);
   }
   return t;
}

// CSR, T lanes per row; result valid in the group's lane 0
template <int T>
__device__ __forceinline__ double rowsum_csr(const int *__restrict__ rp, const int *__restrict__ ci,
                                             const double *__restrict__ va, int row, int sub,
                                             const double *__restrict__ x)
{
   const int beg = rp[row], end = rp[row + 1];
   double t = 0.0;
   int k = beg + sub;
   for (; k + T < end; k += 2 * T)
   {
      int    c0 = ld_stream(ci + k), c1 = ld_stream(ci + k + T);
      double v0 = ld_stream(va + k), v1 = ld_stream(va + k + T);
      double x0 = __ldg(x + c0), x1 = __ldg(x + c1);
      t = fma(v0, x0, t); t = fma(v1, x1, t);
   }
   if (k < end) { t = fma(ld_stream(va + k), __ldg(x + ld_stream(ci + k)), t); }
#pragma unroll
   for (int o = T >> 1; o > 0; o >>= 1) { t += __shfl_down_sync(0xffffffffu, t, o, T); }
   return t;
}

// ---------------------------------------------------------------- kernels
// All rows of the diag block.  T == 0: sliced ELL thread-per-row; T > 0: CSR
// with T lanes per row.  SKIP_BND: rows that own offd entries are left to the
// row-list kernel (they need the halo), so this launch overlaps the exchange.
template <int T, bool SKIP_BND, class Epi>
__global__ void __launch_bounds__(ROWK_BLOCK)
rows_kernel(MatView D, const int *__restrict__ offd_rp, const double *__restrict__ x, Epi epi, RedCtx rc)
{
   double red[Epi::NRED > 0 ? Epi::NRED : 1];
#pragma unroll
   for (int k = 0; k < (Epi::NRED > 0 ? Epi::NRED : 1); k++) { red[k] = 0.0; }

   if (T == 0)
   {
      const int lane = threadIdx.x & 31;
      const int n_slices = (D.n_rows + 31) >> 5;
      const int wpb = ROWK_BLOCK / 32;
      for (int slice = blockIdx.x * wpb + (threadIdx.x >> 5); slice < n_slices; slice += gridDim.x * wpb)
      {
         const int row = (slice << 5) + lane;
         double t = rowsum_sell(D, slice, lane, x);
         if (row < D.n_rows)
         {
            if (!SKIP_BND || offd_rp[row + 1] == offd_rp[row]) { epi(row, t, red); }
         }
      }
   }
   else
   {
      constexpr int TT = (T == 0 ? 1 : T);
      const int sub = threadIdx.x & (TT - 1);
      const int gpb = ROWK_BLOCK / TT;
      // pad the loop so every lane of a warp takes part in the shuffles
      const int n_pad = (D.n_rows + (32 / TT) - 1) / (32 / TT) * (32 / TT);
      for (int row = blockIdx.x * gpb + threadIdx.x / TT; row < n_pad; row += gridDim.x * gpb)
      {
         const bool live = row < D.n_rows;
         double t = rowsum_csr<TT>(D.rp, D.ci, D.va, live ? row : D.n_rows - 1, sub, x);
         if (live && sub == 0)
         {
            if (!SKIP_BND || offd_rp[row + 1] == offd_rp[row]) { epi(row, t, red); }
         }
      }
   }
   grid_reduce<Epi::NRED>(red, rc);
}

// Rows from a list (the boundary rows): diag CSR + offd CSR, 4 lanes per row.
template <class Epi>
__global__ void __launch_bounds__(ROWK_BLOCK)
rowlist_kernel(MatView D, MatView O, const int *__restrict__ rows, int n_list,
               const double *__restrict__ x, const double *__restrict__ xg, Epi epi, RedCtx rc)
{
   double red[Epi::NRED > 0 ? Epi::NRED : 1];
#pragma unroll
   for (int k = 0; k < (Epi::NRED > 0 ? Epi::NRED : 1); k++) { red[k] = 0.0; }
   constexpr int TT = 4;
   const int sub = threadIdx.x & (TT - 1);
   const int gpb = ROWK_BLOCK / TT;
   const int n_pad = (n_list + 7) / 8 * 8;
   for (int idx = blockIdx.x * gpb + threadIdx.x / TT; idx < n_pad; idx += gridDim.x * gpb)
   {
      const bool live = idx < n_list;
      const int row = rows[live ? idx : n_list - 1];
      double t = rowsum_csr<TT>(D.rp, D.ci, D.va, row, sub, x);
      t += rowsum_csr<TT>(O.rp, O.ci, O.va, row, sub, xg);
      if (live && sub == 0) { epi(row, t, red); }
   }
   grid_reduce<Epi::NRED>(red, rc);
}
