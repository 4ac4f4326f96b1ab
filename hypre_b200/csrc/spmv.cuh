// spmv.cuh -- the fused row kernels.  Every solve-phase operator that touches a
// sparse matrix is "row sum t_i = sum_j a_ij x_j, then an epilogue on (i, t_i)":
// SpMV, residual, l1-Jacobi, the Chebyshev steps, two-stage GS, restriction
// (with the stored transpose) and prolongation.  The matrix is streamed once,
// the vectors the epilogue needs are streamed once, and up to B200_MAX_RED dot
// products are reduced in the same launch.
#pragma once
#include "common.cuh"
#include "p2p.cuh"

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
   const int *stop;       // device word: non-zero -> the kernel returns immediately (device-side PCG
                          // convergence: iterations queued ahead of the host become no-ops)
};
#define RETURN_IF_STOPPED(rc) do { if ((rc).stop != nullptr && *(rc).stop != 0) { return; } } while (0)

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
// The gathered operand x_j is obtained through the epilogue functor: xval(x, j).  Plain operators
// read x[j]; a fused operator may compute it from other vectors on the fly (Chebyshev with a zero
// initial guess: x_j = ds_j * (c * (ds_j * f_j)), so the vectors r and tmp are never materialised).
// An epilogue may carry per-thread state prepared once at kernel start (device-resident scalars read into
// registers): optional member init(), called on the kernel's own copy of the functor.
template <class E> __device__ __forceinline__ auto epi_init(E &e, int) -> decltype(e.init(), void()) { e.init(); }
template <class E> __device__ __forceinline__ void epi_init(E &, long) {}
// COH = true (the persistent cycle kernel, tail.cuh): the vector may have been written earlier in the SAME kernel by
// another SM, so it is read from L2 instead of through the non-coherent read-only path.
template <bool COH> __device__ __forceinline__ double ld_gather(const double *p) { return COH ? __ldcg(p) : __ldg(p); }
#define EPI_PLAIN_X template <bool COH> __device__ __forceinline__ double xval(const double *__restrict__ x, int j) const { return ld_gather<COH>(x + j); }
struct XPlain { EPI_PLAIN_X };
// sliced-ELL, one thread per row (lane = row within the 32-row slice).
// Memory latency bounds this loop (ncu: >90 % of the stall samples are long-scoreboard) and
// occupancy hides it: entries are taken in chunks of 4 with a guarded (warp-uniform) tail, which
// keeps the kernel at 32 registers = 8 resident blocks per SM.  A chunk of 8 (two round trips for a
// 7-point row) was measured: 54 registers, half the occupancy, 3-10 % slower on 7-pt and 27-pt.
// The fma chain keeps the storage order.
template <class XF>
__device__ __forceinline__ double rowsum_sell(const MatView &D, int slice, int lane,
                                              const double *__restrict__ x, const XF &xf)
{
   const long long off = D.sl_off[slice];
   const int width = (int)((D.sl_off[slice + 1] - off) >> 5);
   const int    *ci = D.sl_ci + off + lane;
   const double *va = D.sl_va + off + lane;
   double t = 0.0;
   for (int k0 = 0; k0 < width; k0 += 4)
   {
      int    c[4];
      double v[4], xv[4];
      const int m = width - k0;          // entries left (uniform)
#pragma unroll
      for (int j = 0; j < 4; j++) { c[j] = (j < m) ? ld_stream(ci + (k0 + j) * 32) : 0; }
#pragma unroll
      for (int j = 0; j < 4; j++) { v[j] = (j < m) ? ld_stream(va + (k0 + j) * 32) : 0.0; }
#pragma unroll
      for (int j = 0; j < 4; j++) { xv[j] = (j < m) ? xf.template xval<false>(x, c[j]) : 0.0; }
#pragma unroll
      for (int j = 0; j < 4; j++) { if (j < m) { t = fma(v[j], xv[j], t); } }
   }
   return t;
}

// Compressed sliced ELL (DMat::cfmt = CFMT_VAL8 / VAL16 / VAL64).  HBM bandwidth bounds these kernels, so
// the matrix is stored in fewer bytes, losslessly: the column of (slice, entry k, lane) is base[slice][k] + a
// 16-bit offset (the 32 rows of a slice are neighbours, so their k-th columns are, too; a (slice, k) whose
// columns span more than 65534 keeps 32-bit columns in sl_ci: base == CFMT_ESC), and a block with at most
// 256 / 65536 distinct values (stencils, and the Galerkin products of stencils) stores 8- / 16-bit codes
// into a dictionary that stays in L1/L2.  12 bytes per entry become 3 or 4 (10 without a dictionary).
// Entries are stored in chunks of 4 per lane (one 8-byte offset load, one 4/8-byte code load and one
// 16-byte broadcast base load per chunk); slice widths are multiples of 4.  Arithmetic and summation order
// are those of the CSR row.
__device__ __forceinline__ unsigned long long ld_stream_u64(const void *p)
{
   unsigned long long v;
   asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(v) : "l"(p));
   return v;
}
__device__ __forceinline__ unsigned ld_stream_u32(const void *p)
{
   unsigned v;
   asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
   return v;
}

// One slice = entries [off, off + 128 n_chunks).  Two forms of the same storage (DMat::zpad):
//   rowsum_sellc    padding lanes carry the offset CFMT_PAD and are skipped (no gather, no fma)
//   rowsum_sellc_z  no markers: a padding lane gathers a column of its OWN row and multiplies it by a
//                   dictionary entry 0.0 (what rowsum_sell does in the plain layout), so the chunk has no
//                   predicates; u8 dictionaries are read from shared memory.  Row sums are those of
//                   rowsum_sellc except that a sum of exactly -0.0 becomes +0.0.
template <int VF, class XF>
__device__ __forceinline__ double rowsum_sellc(const MatView &D, long long off, int n_chunks, int lane,
                                               const double *__restrict__ x, const XF &xf)
{
   const unsigned short *ip = D.c_idx + off + lane * 4;
   const int4 *bp = reinterpret_cast<const int4 *>(D.c_base + (off >> 5));
   double t = 0.0;
   for (int c = 0; c < n_chunks; c++)
   {
      const unsigned long long id = ld_stream_u64(ip + c * 128);
      const int4 b4 = __ldg(bp + c);
      unsigned long long code = 0;
      if (VF == CFMT_VAL8)  { code = ld_stream_u32((const unsigned char *)D.c_code + off + c * 128 + lane * 4); }
      if (VF == CFMT_VAL16) { code = ld_stream_u64((const unsigned short *)D.c_code + off + c * 128 + lane * 4); }
      const int b[4] = { b4.x, b4.y, b4.z, b4.w };
      int    col[4];
      double v[4], xv[4];
      bool   live[4];
#pragma unroll
      for (int j = 0; j < 4; j++)
      {
         const unsigned o = (unsigned)(id >> (16 * j)) & 0xFFFFu;
         live[j] = (o != CFMT_PAD);
         col[j] = b[j] + (int)o;
         // warp-uniform and rare: this (slice, entry) keeps 32-bit columns
         if (b[j] == CFMT_ESC) { col[j] = ld_stream(D.sl_ci + off + (c * 4 + j) * 32 + lane); }
      }
#pragma unroll
      for (int j = 0; j < 4; j++)
      {
         if (VF == CFMT_VAL8)       { v[j] = __ldg(D.c_dict + ((unsigned)(code >> (8 * j)) & 0xFFu)); }
         else if (VF == CFMT_VAL16) { v[j] = __ldg(D.c_dict + ((unsigned)(code >> (16 * j)) & 0xFFFFu)); }
         else { v[j] = live[j] ? ld_stream(D.sl_va + off + (c * 4 + j) * 32 + lane) : 0.0; }
      }
#pragma unroll
      for (int j = 0; j < 4; j++) { xv[j] = live[j] ? xf.template xval<false>(x, col[j]) : 0.0; }
#pragma unroll
      for (int j = 0; j < 4; j++) { if (live[j]) { t = fma(v[j], xv[j], t); } }
   }
   return t;
}

template <int VF, class XF>
__device__ __forceinline__ double rowsum_sellc_z(const MatView &D, long long off, int n_chunks, int lane,
                                                 const double *__restrict__ x, const XF &xf,
                                                 const double *s_dict)
{
   const unsigned short *ip = D.c_idx + off + lane * 4;
   const int4 *bp = reinterpret_cast<const int4 *>(D.c_base + (off >> 5));
   double t = 0.0;
#pragma unroll 1        // unrolled by two it spills at 32 registers
   for (int c = 0; c < n_chunks; c++)
   {
      const unsigned long long id = ld_stream_u64(ip + c * 128);
      unsigned long long code = 0;
      if (VF == CFMT_VAL8)  { code = ld_stream_u32((const unsigned char *)D.c_code + off + c * 128 + lane * 4); }
      if (VF == CFMT_VAL16) { code = ld_stream_u64((const unsigned short *)D.c_code + off + c * 128 + lane * 4); }
      const int4 b4 = __ldg(bp + c);
      double v[4], xv[4];
      if (VF == CFMT_VAL64)
      {
#pragma unroll
         for (int j = 0; j < 4; j++) { v[j] = ld_stream(D.sl_va + off + (c * 4 + j) * 32 + lane); }
      }
      int col[4] = { b4.x + (int)((unsigned)id & 0xFFFFu), b4.y + (int)((unsigned)(id >> 16) & 0xFFFFu),
                     b4.z + (int)((unsigned)(id >> 32) & 0xFFFFu), b4.w + (int)(unsigned)(id >> 48) };
      if ((b4.x | b4.y | b4.z | b4.w) < 0)       // CFMT_ESC has the sign bit: some entry keeps 32-bit columns
      {
         const int *e = D.sl_ci + off + (c * 4) * 32 + lane;
         if (b4.x < 0) { col[0] = ld_stream(e); }
         if (b4.y < 0) { col[1] = ld_stream(e + 32); }
         if (b4.z < 0) { col[2] = ld_stream(e + 64); }
         if (b4.w < 0) { col[3] = ld_stream(e + 96); }
      }
#pragma unroll
      for (int j = 0; j < 4; j++) { xv[j] = xf.template xval<false>(x, col[j]); }
#pragma unroll
      for (int j = 0; j < 4; j++)
      {
         if (VF == CFMT_VAL8)  { v[j] = s_dict[(unsigned)(code >> (8 * j)) & 0xFFu]; }
         if (VF == CFMT_VAL16) { v[j] = __ldg(D.c_dict + ((unsigned)(code >> (16 * j)) & 0xFFFFu)); }
         t = fma(v[j], xv[j], t);
      }
   }
   return t;
}

// Pattern storage (DMat::cfmt == CFMT_PAT8): stencil operators and their first Galerkin products have a
// handful of distinct (column - row, value) pairs ("patterns": 7 for the 7-pt Laplacian, 151 on level 1 of
// its HMIS hierarchy), so an entry is ONE byte -- the index of its pattern in a dictionary of <= 254
// (value, delta) pairs kept in shared memory -- and where all 32 rows of a slice have the same pattern in
// their k-th entry (interior of a stencil) not even that: the byte sits once in the slice header.
// Slice s owns a header of D.aux x 16 bytes at a FIXED stride (no offset array, so the header of the warp's
// next slice is requested while this one is processed and the chain header -> gather loses a link): word 0 =
// number of header words | first lane-word group << 8, then one header word per 4 entries (a pattern code
// per entry, PAT_LANE = "differs between lanes").  For every header word that contains PAT_LANE the lane
// stream (c_idx slot) holds 32 lane words (the lane's codes at the byte positions of the marked entries).
// Padding entries are patterns (delta of the row's own last entry, 0.0): no predicate anywhere.  Level 0 of a
// stencil problem then streams little more than x and y.
constexpr unsigned PAT_LANE = 0xFFu;

// one header word = 4 entries (entries past the longest row of the slice are padding patterns like any other)
template <class XF>
__device__ __forceinline__ double pat_word(unsigned h, const unsigned *&lp, int row, double t,
                                           const double *__restrict__ x, const XF &xf, const double2 *s_pat)
{
   const unsigned m = __vcmpeq4(h, 0xFFFFFFFFu);          // 0xFF in the bytes marked PAT_LANE (warp-uniform)
   if (m) { const unsigned lw = ld_stream_u32(lp); lp += 32; h = (lw & m) | (h & ~m); }
   double v[4], xv[4];
#pragma unroll
   for (int j = 0; j < 4; j++)
   {
      const double2 pd = s_pat[(h >> (8 * j)) & 0xFFu];
      v[j] = pd.x;
      xv[j] = xf.template xval<false>(x, row + __double2loint(pd.y));
   }
#pragma unroll
   for (int j = 0; j < 4; j++) { t = fma(v[j], xv[j], t); }
   return t;
}

// h0 = the first 16 bytes of the slice header (already in registers), hdr = the header in memory.
// word 0: bits 0-7 header words, 8-9 entries of the last word - 1 (the host decoder stops there; the padding
// entries behind them multiply 0.0 with a column of the row itself), 10-31 first lane-word group.
// (Measured and dropped, profiles/r02_fmt_probe_*.jsonl: the dictionary as a kernel parameter read through
// indexed LDC -- twice as slow as the shared-memory lookups; a grid of one wave of resident blocks -- 1.5x slower
// than four waves; 40 / 48 registers per thread -- slower than 32; the decoded header of lane-uniform slices cached
// in registers from one slice to the next at 64 registers / 4 blocks per SM -- 1.8x slower; two header words per trip
// with all eight gathers issued before the first fma at 40 registers / 6 blocks -- 12 % slower.  The loop is bound by
// load latency and occupancy is what hides it: every variant that trades resident warps for work per warp lost.)
template <class XF>
__device__ __forceinline__ double rowsum_pat(const MatView &D, const uint4 &h0, const uint4 *hdr, int lane, int row,
                                             const double *__restrict__ x, const XF &xf, const double2 *s_pat)
{
   const unsigned nw = h0.x & 0xFFu;
   const unsigned *lp = reinterpret_cast<const unsigned *>(D.c_idx) + (size_t)(h0.x >> 10) * 32 + lane;
   double t = 0.0;
   if (nw > 0) { t = pat_word(h0.y, lp, row, t, x, xf, s_pat); }
   if (nw > 1) { t = pat_word(h0.z, lp, row, t, x, xf, s_pat); }
   if (nw > 2) { t = pat_word(h0.w, lp, row, t, x, xf, s_pat); }
   const unsigned *hw = reinterpret_cast<const unsigned *>(hdr) + 1;
   for (unsigned i = 3; i < nw; i++) { t = pat_word(__ldg(hw + i), lp, row, t, x, xf, s_pat); }
   return t;
}

// CG: gather x with ld.global.cg (L2 only) -- for ghost vectors that a peer GPU wrote during this
// kernel's lifetime, which must not be served from a stale L1 / read-only cache line
template <bool CG>
__device__ __forceinline__ double ld_x(const double *p) { return CG ? __ldcg(p) : __ldg(p); }

template <int T, bool CG, class XF>
__device__ __forceinline__ double rowsum_csr(const int *__restrict__ rp, const int *__restrict__ ci,
                                             const double *__restrict__ va, int row, int sub,
                                             const double *__restrict__ x, const XF &xf)
{
   const int beg = rp[row], end = rp[row + 1];
   double t = 0.0;
   int k = beg + sub;
   for (; k + T < end; k += 2 * T)
   {
      int    c0 = ld_stream(ci + k), c1 = ld_stream(ci + k + T);
      double v0 = ld_stream(va + k), v1 = ld_stream(va + k + T);
      double x0 = CG ? __ldcg(x + c0) : xf.template xval<false>(x, c0), x1 = CG ? __ldcg(x + c1) : xf.template xval<false>(x, c1);
      t = fma(v0, x0, t); t = fma(v1, x1, t);
   }
   if (k < end)
   {
      const int c0 = ld_stream(ci + k);
      t = fma(ld_stream(va + k), CG ? __ldcg(x + c0) : xf.template xval<false>(x, c0), t);
   }
#pragma unroll
   for (int o = T >> 1; o > 0; o >>= 1) { t += __shfl_down_sync(0xffffffffu, t, o, T); }
   return t;
}

// ---------------------------------------------------------------- kernels
// What a launch does about the off-process part of the operator:
//   HALO_NONE   single rank, or an operator without ghosts (transposes, strict lower triangle)
//   HALO_SKIP   NCCL path: rows that own offd entries are left to rowlist_kernel (they need the
//               halo), so this launch overlaps the library exchange running on the comm stream
//   HALO_FWD    peer-memory path, y-type operators (hypre_ParCSRMatrixMatvecOutOfPlace,
//               src/parcsr_mv/par_csr_matvec.c:163-218) as ONE kernel: the first blocks pack
//               x[send_map] and store it straight into the neighbours' ghost vectors (NVLink),
//               all blocks then stream their rows; a warp that meets a row with offd entries waits
//               (once) for the neighbours' sequence numbers, adds the offd part from the ghost
//               vector and goes on -- interior rows never wait, the transfer overlaps the math
//   HALO_REV    peer-memory path, A^T x (hypre_ParCSRMatrixMatvecT, par_csr_matvec.c:287-514):
//               the first blocks compute offd^T x and store it into the owners' receive buffers,
//               all blocks stream diag^T; rows that receive contributions wait and add them
//               in a fixed order (unpack lists grouped by row)
enum { HALO_NONE = 0, HALO_SKIP = 1, HALO_FWD = 2, HALO_REV = 3 };

struct HaloArgs {
   MatView     O;        // SKIP: offd (rp only); FWD: offd block; REV: offd^T block (rows = ghost columns)
   P2PChan     ch;       // FWD: forward channel; REV: reverse channel
   const int  *map;      // FWD: send_map_elmts
   const int  *urp;      // REV: [n_rows+1] contributions per local row
   const int  *upos;     // REV: positions in the receive buffer, grouped by row
   int         dbg;      // B200_P2P_DEBUG bits (timing experiments only):
                         // 2 no waiting, 4 no flow control / acknowledgements, 8 no system fences in the put
   int         int_begin, int_end;   // FWD/REV: longest run of 32-row groups that need no remote data
   const unsigned char *grp;   // FWD/REV: [ceil(n_rows/32)] non-zero if a row of that 32-row group needs remote data
   int         chunked;  // thread-per-row layouts: contiguous run of slices per block instead of a grid stride
};

// sender part of HALO_FWD / HALO_REV, executed by blocks [0, n_put_blocks): values go straight into the receivers'
// buffers, each with the sequence number of this exchange (p2p.cuh: flag-in-data) -- no fence, no flag
template <int HALO, class XF>
__device__ __forceinline__ void halo_put_phase(const HaloArgs &h, const double *__restrict__ x, int n_put_blocks,
                                               const XF &xf)
{
   __shared__ u64  s_seq;
   __shared__ bool s_last_p;
   const P2PChan &ch = h.ch;
   if (threadIdx.x == 0)
   {
      const u64 seq = ld_relaxed_gpu(ch.seq_put) + 1;
      // flow control: buffer (seq & 1) was used by exchange seq-2; receivers that do not send to me acknowledge
      if (seq > 2 && !(h.dbg & 4))
      {
         for (int i = 0; i < ch.n_out; i++) { if (ch.out_wait_ack[i]) { p2p_spin_ge(ch.out_ack + i, seq - 2, ch.err); } }
      }
      s_seq = seq;
   }
   __syncthreads();
   const u64 seq = s_seq;
   const unsigned flag = (unsigned)seq;
   LLVal *const *dst = ch.out_dst[(int)(seq & 1)];
   int seg = 0;
   for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < ch.n_out_total; i += n_put_blocks * blockDim.x)
   {
      while (i >= ch.out_starts[seg + 1]) { seg++; }
      double v;
      if (HALO == HALO_FWD) { v = xf.template xval<false>(x, h.map[i]); }   // pack: x_buf[i] = x_local[send_map_elmts[i]]
      else
      {
         // row i of offd^T: contribution of my rows to ghost column i
         v = 0.0;
         for (int k = h.O.rp[i]; k < h.O.rp[i + 1]; k++) { v = fma(h.O.va[k], x[h.O.ci[k]], v); }
      }
      ll_store(dst[seg] + (i - ch.out_starts[seg]), v, flag);
   }
   // the last put block advances the local sequence counter (every put block has read it by then)
   __syncthreads();
   if (n_put_blocks == 1) { s_last_p = true; }
   else
   {
      if (threadIdx.x == 0) { s_last_p = (atomicAdd(ch.ticket, 1u) == (unsigned)n_put_blocks - 1u); }
      __syncthreads();
   }
   if (s_last_p && threadIdx.x == 0) { *ch.seq_put = seq; ch.ticket[0] = 0u; }
}

// receiver part: called by a converged warp when one of its rows needs remote data; fixes the buffer and the
// sequence number this exchange's values carry (once per warp).  The waiting itself is per element (ll_read).
__device__ __forceinline__ const LLVal *halo_warp_wait(const P2PChan &ch, const LLVal *cur, unsigned &flag)
{
   if (cur != nullptr) { return cur; }
   unsigned f = 0;
   if ((threadIdx.x & 31) == 0) { f = (unsigned)(ld_relaxed_gpu(ch.seq_wait) + 1); }
   f = __shfl_sync(0xffffffffu, f, 0);
   flag = f;
   return ch.in_buf[(int)(f & 1u)];
}

// CSR rows of the offd block over the receive buffer (T lanes per row, all lanes of the warp take part)
template <int T>
__device__ __forceinline__ double rowsum_csr_ll(const int *__restrict__ rp, const int *__restrict__ ci,
                                                const double *__restrict__ va, int row, int sub,
                                                const LLVal *xg, unsigned flag, int *err)
{
   const int beg = rp[row], end = rp[row + 1];
   double t = 0.0;
   for (int k = beg + sub; k < end; k += T) { t = fma(ld_stream(va + k), ll_read(xg + ld_stream(ci + k), flag, err), t); }
#pragma unroll
   for (int o = T >> 1; o > 0; o >>= 1) { t += __shfl_down_sync(0xffffffffu, t, o, T); }
   return t;
}

// Row groups [g_begin, g_end) (a group = 32 consecutive rows = one slice of the sliced-ELL layouts) of the
// diag block.  T == 0: sliced ELL thread-per-row; T < 0: compressed thread-per-row storage, format
// (-T) & 7 (CFMT_*), variant (-T) >> 3 (VAL*: 1 = storage without padding markers, rowsum_sellc_z; PAT8:
// 1 = the instantiation with more registers per thread); T > 0: CSR with T lanes per row.
// HALO as above; HALO_NONE / HALO_SKIP never touch the channel.
// Thread-per-row mapping: h.chunked != 0 gives every block ONE contiguous run of slices (its warps walk it
// 8 slices at a time), so that the gathers of neighbouring rows -- the +-nx lines of a stencil -- meet in the
// SM's L1 instead of each costing an L2 read; 0 = grid stride.
template <int T, int HALO, class Epi>
__device__ __forceinline__ void rows_range(const MatView &D, const HaloArgs &h, const double *__restrict__ x,
                                           const Epi &epi, double *red, int g_begin, int g_end, const LLVal *&xg,
                                           unsigned &xflag, int vblock, const double2 *s_tab = nullptr)
{
   // vblock: the block's position in this range's grid (a permutation of blockIdx.x), so that
   // short ranges do not all land on the low-numbered blocks, which also carry the put phase
   if (T <= 0)
   {
      constexpr int FMT  = T < 0 ? ((-T) & 7) : 0;
      constexpr bool ZPAD = T < 0 && FMT != CFMT_PAT8 && ((-T) >> 3) == 1;
      const int lane = threadIdx.x & 31;
      const int wpb = ROWK_BLOCK / 32;
      int slice, s_end, step;
      if (h.chunked)
      {
         const int per_block = ((g_end - g_begin + (int)gridDim.x * wpb - 1) / ((int)gridDim.x * wpb)) * wpb;
         slice = g_begin + vblock * per_block;
         s_end = slice + per_block < g_end ? slice + per_block : g_end;
         slice += (threadIdx.x >> 5);
         step = wpb;
      }
      else
      {
         slice = g_begin + vblock * wpb + (threadIdx.x >> 5);
         s_end = g_end;
         step = gridDim.x * wpb;
      }
      const uint4 *hdr = reinterpret_cast<const uint4 *>(D.c_code);
      uint4 h_cur = make_uint4(0u, 0u, 0u, 0u);
      if (FMT == CFMT_PAT8 && slice < s_end) { h_cur = __ldg(hdr + (size_t)slice * D.aux); }
      for (; slice < s_end; slice += step)
      {
         int row = (slice << 5) + lane;
         if (D.perm != nullptr) { row = ((slice >> 3) << 8) + D.perm[row]; }   // sorted inside 256-row windows
         const bool live = row < D.n_rows;
         double t;
         if (T == 0) { t = rowsum_sell(D, slice, lane, x, epi); }
         else if (FMT == CFMT_PAT8)
         {
            // the header of the warp's next slice travels while this one is processed
            uint4 h_next = make_uint4(0u, 0u, 0u, 0u);
            if (slice + step < s_end) { h_next = __ldg(hdr + (size_t)(slice + step) * D.aux); }
            t = rowsum_pat(D, h_cur, hdr + (size_t)slice * D.aux, lane, live ? row : D.n_rows - 1, x, epi, s_tab);
            h_cur = h_next;
         }
         else
         {
            const long long off = D.sl_off[slice];
            const int nch = (int)((D.sl_off[slice + 1] - off) >> 7);
            t = ZPAD ? rowsum_sellc_z<FMT>(D, off, nch, lane, x, epi, reinterpret_cast<const double *>(s_tab))
                     : rowsum_sellc<FMT>(D, off, nch, lane, x, epi);
         }
         bool skip = false;
         if (HALO == HALO_SKIP) { skip = live && h.O.rp[row + 1] != h.O.rp[row]; }
         if (HALO == HALO_FWD && h.grp[slice])
         {
            const int ob = live ? h.O.rp[row] : 0, oe = live ? h.O.rp[row + 1] : 0;
            xg = halo_warp_wait(h.ch, xg, xflag);
            for (int k = ob; k < oe; k++) { t = fma(h.O.va[k], ll_read(xg + h.O.ci[k], xflag, h.ch.err), t); }
         }
         if (HALO == HALO_REV && h.grp[slice])
         {
            const int ub = live ? h.urp[row] : 0, ue = live ? h.urp[row + 1] : 0;
            xg = halo_warp_wait(h.ch, xg, xflag);
            double e = 0.0;
            for (int k = ub; k < ue; k++) { e += ll_read(xg + h.upos[k], xflag, h.ch.err); }
            t += e;
         }
         if (live && !skip) { epi(row, t, red); }
      }
   }
   else
   {
      constexpr int TT = (T <= 0 ? 1 : T);
      const int sub = threadIdx.x & (TT - 1);
      const int gpb = ROWK_BLOCK / TT;
      // whole warps: a warp's 32/TT rows lie in one group, so the group flag is warp-uniform and every
      // lane takes part in the shuffles
      const int r_end = g_end << 5;
      for (int row = (g_begin << 5) + vblock * gpb + threadIdx.x / TT; row < r_end; row += gridDim.x * gpb)
      {
         const bool live = row < D.n_rows;
         const int r = live ? row : D.n_rows - 1;
         double t = rowsum_csr<TT, false>(D.rp, D.ci, D.va, r, sub, x, epi);
         bool skip = false;
         if (HALO == HALO_SKIP) { skip = h.O.rp[r + 1] != h.O.rp[r]; }
         if (HALO == HALO_FWD && h.grp[row >> 5])
         {
            xg = halo_warp_wait(h.ch, xg, xflag);
            // all lanes take part (rows without offd entries have an empty range)
            t += rowsum_csr_ll<TT>(h.O.rp, h.O.ci, h.O.va, r, sub, xg, xflag, h.ch.err);
         }
         if (HALO == HALO_REV && h.grp[row >> 5])
         {
            const int ub = live ? h.urp[r] : 0, ue = live ? h.urp[r + 1] : 0;
            xg = halo_warp_wait(h.ch, xg, xflag);
            if (sub == 0)
            {
               double e = 0.0;
               for (int k = ub; k < ue; k++) { e += ll_read(xg + h.upos[k], xflag, h.ch.err); }
               t += e;
            }
         }
         if (live && sub == 0 && !skip) { epi(row, t, red); }
      }
   }
}

// resident blocks per SM: 8 x 256 threads at 32 registers hide the DRAM latency of the streaming kernels;
// the second PAT8 instantiation trades occupancy for registers (the four gathers of a header word in flight)
// Epilogues whose gathered operand is composite (two vectors read per entry: EpiChebyZero, EpiMatvecDotDir) spill at
// 32 (and 40) registers on the thread-per-row kernels; they run at 48 registers / 5 blocks per SM (specialised in ops.cu).
template <class Epi> struct EpiHeavy { static constexpr bool value = false; };

template <int T, class Epi> struct RowsKernelCfg {
   static constexpr int FMT = T < 0 ? ((-T) & 7) : 0;
   static constexpr int VAR = T < 0 ? ((-T) >> 3) : 0;
   static constexpr bool PAT = FMT == CFMT_PAT8;
   static constexpr bool SMEM_TAB = PAT || (VAR == 1 && FMT == CFMT_VAL8);
   static constexpr int MIN_BLOCKS = (T <= 0 && EpiHeavy<Epi>::value) ? 5 : 8;
};

// All rows of the diag block.
template <int T, int HALO, class Epi>
__global__ void __launch_bounds__(ROWK_BLOCK, RowsKernelCfg<T, Epi>::MIN_BLOCKS)
rows_kernel(MatView D, HaloArgs h, const double *__restrict__ x, Epi epi_arg, RedCtx rc)
{
   RETURN_IF_STOPPED(rc);
   Epi epi = epi_arg;
   epi_init(epi, 0);
   double red[Epi::NRED > 0 ? Epi::NRED : 1];
#pragma unroll
   for (int k = 0; k < (Epi::NRED > 0 ? Epi::NRED : 1); k++) { red[k] = 0.0; }

   constexpr bool P2P = (HALO == HALO_FWD || HALO == HALO_REV);
   using Cfg = RowsKernelCfg<T, Epi>;
   // pattern dictionary (value, delta) / u8 value dictionary, staged once per block
   __shared__ double2 s_tab[Cfg::SMEM_TAB ? 256 : 1];
   if (Cfg::SMEM_TAB)
   {
      if (Cfg::PAT) { if ((int)threadIdx.x < D.n_dict) { s_tab[threadIdx.x] = reinterpret_cast<const double2 *>(D.c_dict)[threadIdx.x]; } }
      else if ((int)threadIdx.x < D.n_dict) { reinterpret_cast<double *>(s_tab)[threadIdx.x] = D.c_dict[threadIdx.x]; }
      __syncthreads();
   }
   const int n_groups = (D.n_rows + 31) >> 5;
   const LLVal *xg = nullptr;     // receive buffer of this exchange once this warp has looked it up
   unsigned xflag = 0;            // sequence number its values carry
   if (!P2P) { rows_range<T, HALO, Epi>(D, h, x, epi, red, 0, n_groups, xg, xflag, blockIdx.x, s_tab); }
   else
   {
      if (h.ch.n_out > 0)
      {
         // the put blocks must all be co-resident with the blocks that spin on the peers' flags (the peer is
         // in the same state): at most one wave of the smallest residency, the pack is grid-strided
         int n_put_blocks = (h.ch.n_out_total + ROWK_BLOCK - 1) / ROWK_BLOCK;
         if (n_put_blocks > B200_NUM_SMS * 4) { n_put_blocks = B200_NUM_SMS * 4; }
         if (n_put_blocks > (int)gridDim.x) { n_put_blocks = gridDim.x; }
         if ((int)blockIdx.x < n_put_blocks) { halo_put_phase<HALO>(h, x, n_put_blocks, epi); }
      }
      // interior groups [int_begin, int_end) first, with exactly the single-rank code; the groups that
      // need remote data (both ends of a 1-D block partition) last, when the halo has long arrived
      const int G = gridDim.x, b = blockIdx.x;
      rows_range<T, HALO_NONE, Epi>(D, h, x, epi, red, h.int_begin, h.int_end, xg, xflag, b, s_tab);
      rows_range<T, HALO, Epi>(D, h, x, epi, red, 0, h.int_begin, xg, xflag, G - 1 - b, s_tab);
      rows_range<T, HALO, Epi>(D, h, x, epi, red, h.int_end, n_groups, xg, xflag, (b + G / 2) % G, s_tab);
   }
   grid_reduce<Epi::NRED>(red, rc);
   if (P2P && h.ch.n_in > 0 && !(h.dbg & 4)) { p2p_consumer_done(h.ch); }
}

// Rows from a list (the boundary rows of the NCCL path): diag CSR + offd CSR, 4 lanes per row.
template <class Epi>
__global__ void __launch_bounds__(ROWK_BLOCK)
rowlist_kernel(MatView D, MatView O, const int *__restrict__ rows, int n_list,
               const double *__restrict__ x, const double *__restrict__ xg, Epi epi_arg, RedCtx rc)
{
   RETURN_IF_STOPPED(rc);
   Epi epi = epi_arg;
   epi_init(epi, 0);
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
      double t = rowsum_csr<TT, false>(D.rp, D.ci, D.va, row, sub, x, epi);
      t += rowsum_csr<TT, false>(O.rp, O.ci, O.va, row, sub, xg, XPlain());    // ghosts hold finished values
      if (live && sub == 0) { epi(row, t, red); }
   }
   grid_reduce<Epi::NRED>(red, rc);
}
