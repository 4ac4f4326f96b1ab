// noalloc.cu -- an inner product that allocates nothing (no counterpart in the reference: added
// for callers that cannot call cudaMalloc -- code under CUDA graph capture, tight serving
// loops, harnesses that forbid it).
//
// b200_vec_dot (ctx.cu / vecops.cu) reduces through scratch memory the context allocated at creation
// (ctx->d_partials, ctx->d_counter).  The variant here keeps that scratch in a file-scope device array of a
// stated maximum size instead, so that neither this call nor any set-up it needs touches an allocator:
//
//   * g_dot_scratch: B200_NOALLOC_DOT_MAX_BLOCKS block partials (128 KB) + one ticket word, held by every GPU
//     that loads the library until the process exits;
//   * in host code the array's NAME is only a host-side stand-in: the device address is looked up once per
//     device, with that device current, by cudaGetSymbolAddress (checked), and passed to the kernel;
//   * every call shares the array and it keeps the last call's contents: the ticket word is cleared on the
//     caller's stream before each launch, the partials are fully overwritten before they are read.  Calls
//     that OVERLAP on the GPU (two streams, two contexts on one device) race on it -- serialise them.
#include "common.cuh"

constexpr int NOALLOC_DOT_BLOCK = 256;
constexpr int NOALLOC_DOT_PER_THREAD = 16;
constexpr int NOALLOC_DOT_PER_BLOCK = NOALLOC_DOT_BLOCK * NOALLOC_DOT_PER_THREAD;          // 4096 elements per block

struct NoallocDotScratch {
   double   partial[B200_NOALLOC_DOT_MAX_BLOCKS];
   unsigned ticket;
};
static __device__ __align__(256) NoallocDotScratch g_dot_scratch;

static_assert((long long)B200_NOALLOC_DOT_MAX_BLOCKS * NOALLOC_DOT_PER_BLOCK == (long long)B200_NOALLOC_DOT_MAX_N,
              "B200_NOALLOC_DOT_MAX_N = blocks x elements per block");

// block b reduces elements [4096 b, 4096 (b + 1)) and stores its partial; the last block to finish (ticket) sums
// the partials in index order: deterministic for a given n, result written to the caller's device word
__global__ void __launch_bounds__(NOALLOC_DOT_BLOCK)
noalloc_dot_kernel(int n, const double *__restrict__ x, const double *__restrict__ y, NoallocDotScratch *s,
                   double *__restrict__ out)
{
   __shared__ double s_red[NOALLOC_DOT_BLOCK / 32];
   __shared__ bool   s_last;
   const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
   const long long base = (long long)blockIdx.x * NOALLOC_DOT_PER_BLOCK;
   double v = 0.0;
#pragma unroll 4
   for (int k = 0; k < NOALLOC_DOT_PER_THREAD; k++)
   {
      const long long i = base + (long long)k * NOALLOC_DOT_BLOCK + threadIdx.x;
      if (i < n) { v = fma(x[i], y[i], v); }
   }
#pragma unroll
   for (int o = 16; o > 0; o >>= 1) { v += __shfl_down_sync(0xffffffffu, v, o); }
   if (lane == 0) { s_red[warp] = v; }
   __syncthreads();
   if (threadIdx.x == 0)
   {
      double t = 0.0;
      for (int w = 0; w < NOALLOC_DOT_BLOCK / 32; w++) { t += s_red[w]; }
      s->partial[blockIdx.x] = t;
      __threadfence();
      s_last = (atomicAdd(&s->ticket, 1u) == gridDim.x - 1);
   }
   __syncthreads();
   if (!s_last) { return; }
   __threadfence();
   double t = 0.0;
   for (int b = threadIdx.x; b < (int)gridDim.x; b += NOALLOC_DOT_BLOCK) { t += ((volatile double *)s->partial)[b]; }
#pragma unroll
   for (int o = 16; o > 0; o >>= 1) { t += __shfl_down_sync(0xffffffffu, t, o); }
   __syncthreads();
   if (lane == 0) { s_red[warp] = t; }
   __syncthreads();
   if (threadIdx.x == 0)
   {
      double r = 0.0;
      for (int w = 0; w < NOALLOC_DOT_BLOCK / 32; w++) { r += s_red[w]; }
      *out = r;
      s->ticket = 0u;
   }
}

// device address of g_dot_scratch, per device (looked up once, with that device current)
constexpr int NOALLOC_MAX_DEVICES = 64;
static NoallocDotScratch *g_scratch_addr[NOALLOC_MAX_DEVICES] = {nullptr};

static int noalloc_scratch(int device, NoallocDotScratch **out)
{
   ARG_CHECK(device >= 0 && device < NOALLOC_MAX_DEVICES);
   if (g_scratch_addr[device] == nullptr)
   {
      int cur = -1;
      CUDA_TRY(cudaGetDevice(&cur));
      if (cur != device) { CUDA_TRY(cudaSetDevice(device)); }
      void *p = nullptr;
      CUDA_TRY(cudaGetSymbolAddress(&p, g_dot_scratch));
      if (p == nullptr) { b200_set_error("cudaGetSymbolAddress returned a null address for the dot scratch"); return B200_ERR_CUDA; }
      g_scratch_addr[device] = static_cast<NoallocDotScratch *>(p);
   }
   *out = g_scratch_addr[device];
   return B200_OK;
}

extern "C" int b200_vec_dot_noalloc(b200_ctx *c, int n, const double *x, const double *y, double *d_result)
{
   ARG_CHECK(c != nullptr && n >= 0 && d_result != nullptr && (n == 0 || (x != nullptr && y != nullptr)));
   if (n > B200_NOALLOC_DOT_MAX_N)
   {
      b200_set_error("b200_vec_dot_noalloc: n = %d exceeds the static scratch (B200_NOALLOC_DOT_MAX_N = %d elements); "
                     "use b200_vec_dot", n, (int)B200_NOALLOC_DOT_MAX_N);
      return B200_ERR_ARG;
   }
   CUDA_TRY(cudaSetDevice(c->device));
   NoallocDotScratch *s = nullptr;
   B200_TRY(noalloc_scratch(c->device, &s));
   if (n == 0)
   {
      CUDA_TRY(cudaMemsetAsync(d_result, 0, sizeof(double), c->stream));     // +0.0
      return B200_OK;
   }
   // the array keeps the last call's contents (also those of a call that was aborted half way): the ticket starts at zero
   CUDA_TRY(cudaMemsetAsync(&s->ticket, 0, sizeof(unsigned), c->stream));
   const int grid = (n + NOALLOC_DOT_PER_BLOCK - 1) / NOALLOC_DOT_PER_BLOCK;
   noalloc_dot_kernel<<<grid, NOALLOC_DOT_BLOCK, 0, c->stream>>>(n, x, y, s, d_result);
   c->launches++;
   CHECK_LAUNCH();
   return B200_OK;
}
