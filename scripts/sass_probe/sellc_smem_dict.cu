#include <cstdint>
constexpr int CFMT_ESC = (int)0x80000000;
constexpr unsigned CFMT_PAD = 0xFFFFu;
__device__ __forceinline__ unsigned long long ld_stream_u64(const void *p)
{ unsigned long long v; asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(v) : "l"(p)); return v; }
__device__ __forceinline__ unsigned ld_stream_u32(const void *p)
{ unsigned v; asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p)); return v; }
__device__ __forceinline__ int ld_stream_i(const int *p)
{ int v; asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(v) : "l"(p)); return v; }
struct MV { int n_rows; const long long *sl_off; const unsigned short *c_idx; const int *c_base; const unsigned char *c_code; const double *c_dict; const int *sl_ci; int n_dict; };

__global__ void __launch_bounds__(256, 8) probe(MV D, const double *__restrict__ x, double *__restrict__ y, int n_slices)
{
   __shared__ double s_dict[256];
   if (threadIdx.x < D.n_dict) { s_dict[threadIdx.x] = D.c_dict[threadIdx.x]; }
   __syncthreads();
   const int lane = threadIdx.x & 31;
   for (int slice = blockIdx.x * 8 + (threadIdx.x >> 5); slice < n_slices; slice += gridDim.x * 8)
   {
      const long long off = D.sl_off[slice];
      const int n_chunks = (int)((D.sl_off[slice + 1] - off) >> 7);
      const unsigned short *ip = D.c_idx + off + lane * 4;
      const unsigned char *cp = D.c_code + off + lane * 4;
      const int4 *bp = reinterpret_cast<const int4 *>(D.c_base + (off >> 5));
      double t = 0.0;
      for (int c = 0; c < n_chunks; c++)
      {
         const unsigned long long id = ld_stream_u64(ip + c * 128);
         const unsigned code = ld_stream_u32(cp + c * 128);
         const int4 b4 = __ldg(bp + c);
         int col[4] = { b4.x + (int)((unsigned)id & 0xFFFFu), b4.y + (int)((unsigned)(id >> 16) & 0xFFFFu),
                        b4.z + (int)((unsigned)(id >> 32) & 0xFFFFu), b4.w + (int)(unsigned)(id >> 48) };
         if ((b4.x | b4.y | b4.z | b4.w) < 0)       // some entry of this chunk keeps 32-bit columns (rare, warp-uniform)
         {
            const int *e = D.sl_ci + off + (c * 4) * 32 + lane;
            if (b4.x < 0) { col[0] = ld_stream_i(e); }
            if (b4.y < 0) { col[1] = ld_stream_i(e + 32); }
            if (b4.z < 0) { col[2] = ld_stream_i(e + 64); }
            if (b4.w < 0) { col[3] = ld_stream_i(e + 96); }
         }
         double xv[4];
#pragma unroll
         for (int j = 0; j < 4; j++)
         {
            const bool live = ((unsigned)(id >> (16 * j)) & 0xFFFFu) != CFMT_PAD;
            xv[j] = live ? __ldg(x + col[j]) : 0.0;
         }
#pragma unroll
         for (int j = 0; j < 4; j++)
         {
            const bool live = ((unsigned)(id >> (16 * j)) & 0xFFFFu) != CFMT_PAD;
            const double v = s_dict[(code >> (8 * j)) & 0xFFu];
            if (live) { t = fma(v, xv[j], t); }
         }
      }
      const int row = slice * 32 + lane;
      if (row < D.n_rows) { y[row] = t; }
   }
}
