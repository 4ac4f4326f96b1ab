// ops.cu -- SpMV, residual, smoothers and transfer operators as epilogues of the
// fused row kernels (spmv.cuh).
#include "ops.cuh"
#include "vecops.cuh"

// ---------------------------------------------------------------- epilogues
// Arithmetic of the epilogues is pinned with __dmul_rn / __dadd_rn: separately rounded products and sums, as the
// reference's host code (gcc -O2, x86-64 baseline: no fused multiply-add) evaluates the same expressions, and
// independent of how nvcc would contract a*b+c in one inlining context or another (the stand-alone launch and the
// persistent coarse-level kernel must agree bit for bit).  The row sums themselves are explicit fma chains.
#define MUL(a, b) __dmul_rn((a), (b))
#define ADD(a, b) __dadd_rn((a), (b))
#define SUB(a, b) __dsub_rn((a), (b))
// y = alpha*t + beta*b           hypre_CSRMatrixMatvecOutOfPlaceHost (src/seq_mv/csr_matvec.c:683-847)
struct EpiMatvec {
   static constexpr int NRED = 0;
   EPI_PLAIN_X
   double alpha, beta; const double *b; double *y;
   __device__ void operator()(int r, double t, double *) const
   {
      double v = MUL(alpha, t);
      if (beta != 0.0) { v = ADD(v, MUL(beta, b[r])); }      // y = alpha*(A x) + beta*b, two rounded products (csr_matvec.c)
      y[r] = v;
   }
};
// y = t, red0 += y*x             PCG: s = A p and <s,p> (src/krylov/pcg.c:577-580)
struct EpiMatvecDot {
   static constexpr int NRED = 1;
   EPI_PLAIN_X
   const double *x; double *y;
   __device__ void operator()(int r, double t, double *red) const
   {
      y[r] = t;
      red[0] = fma(t, x[r], red[0]);
   }
};
// PCG: the direction update p = z + beta p (src/krylov/pcg.c:960-971: ScaleVector(beta, p); Axpy(1, s, p)) fused
// into the next s = A p and <s, p> (pcg.c:577-580): the gathered operand is formed on the fly from z and the OLD p,
// the epilogue writes the new p into the other of two buffers (other rows still gather the old one) -- p is streamed
// once per iteration instead of three times and the pass of its own disappears.  beta and the parity of the
// iteration (which buffer is old) are device-resident: the launch is identical from one iteration to the next.
struct EpiMatvecDotDir {
   static constexpr int NRED = 1;
   const double *scal_beta; const int *iter; const double *z; double *pbuf0; double *pbuf1; double *s;
   double beta; const double *p_old; double *p_new;       // set by init()
   __device__ void init()
   {
      beta = *scal_beta;
      const int odd = *iter & 1;
      p_old = odd ? pbuf1 : pbuf0;
      p_new = odd ? pbuf0 : pbuf1;
   }
   template <bool COH> __device__ __forceinline__ double xval(const double *__restrict__, int j) const
   {
      return fma(beta, ld_gather<COH>(p_old + j), ld_gather<COH>(z + j));
   }
   __device__ void operator()(int r, double t, double *red) const
   {
      const double pn = fma(beta, p_old[r], z[r]);
      p_new[r] = pn;
      s[r] = t;
      red[0] = fma(t, pn, red[0]);
   }
};
// u_out = u_in + (w*(f - t)) / l1     hypre_BoomerAMGRelax7Jacobi (src/parcsr_ls/par_relax.c:1177-1254)
// DOT: the sweep is the last operation of a preconditioner application z = C r inside PCG, so gamma = <r, z>
// (src/krylov/pcg.c:737) is reduced here, while z_i is in a register, instead of in a pass of its own.
template <bool DOT>
struct EpiJacobiT {
   static constexpr int NRED = DOT ? 1 : 0;
   EPI_PLAIN_X
   double w; const double *f; const double *l1; const double *u_in; double *u_out; const double *dw;
   const int *marker; int points;     // CF-ordered sweep (relax_order 1): only rows with marker == points move
   __device__ void operator()(int r, double t, double *red) const
   {
      // hypre_ParVectorElmdivpyMarked (seq_mv/vector.c:823-870): u_i += Vtemp_i / l1_i where marker_i == points
      const bool on = marker == nullptr || marker[r] == points;
      const double un = on ? ADD(u_in[r], __ddiv_rn(MUL(w, SUB(f[r], t)), l1[r])) : u_in[r];
      u_out[r] = un;
      if (DOT) { red[0] = fma(dw[r], un, red[0]); }
   }
};
// Chebyshev first step (src/parcsr_ls/par_cheby.c:275-283,318-341):
//   r = ds.*(f - t)  [scale]  |  r = f - t ;  tmp = ds.*(c*r) | c*r
struct EpiChebyFirst {
   static constexpr int NRED = 0;
   EPI_PLAIN_X
   int scale; double c; const double *f; const double *ds; double *rv; double *tmp;
   __device__ void operator()(int r, double t, double *) const
   {
      double res = SUB(f[r], t);
      if (scale) { res = MUL(ds[r], res); }
      rv[r] = res;
      double u1 = MUL(res, c);
      tmp[r] = scale ? MUL(ds[r], u1) : u1;
   }
};
// Chebyshev inner step (par_cheby.c:284-297,344-371): u_i = c_i*r + ds.*t ; then either
//   out = ds.*u_i (next SpMV input)   or   u += ds.*u_i (last step)
template <bool DOT>
struct EpiChebyStepT {
   static constexpr int NRED = DOT ? 1 : 0;
   EPI_PLAIN_X
   int scale; int last; double c; const double *ds; const double *rv; double *out; double *u; const double *dw;
   __device__ void operator()(int r, double t, double *red) const
   {
      double ui = scale ? ADD(MUL(c, rv[r]), MUL(ds[r], t)) : ADD(MUL(c, rv[r]), t);
      double su = scale ? MUL(ds[r], ui) : ui;
      if (last)
      {
         const double un = ADD(u[r], su);
         u[r] = un;
         if (DOT) { red[0] = fma(dw[r], un, red[0]); }      // last step of the last sweep: <r, z> of PCG
      }
      else      { out[r] = su; }
   }
};
// Chebyshev with a zero initial guess (u == 0 on entry: PCG's ClearVector, the cycle's U[l+1] = 0):
// the first step needs no SpMV (A*0 == 0): r = ds.*f, tmp = ds.*(c_k*r).  Instead of writing r and
// tmp in an elementwise pass, the next SpMV computes its gathered operand tmp_j from f_j, ds_j on the
// fly (xval) and recomputes r_i in the epilogue -- the same floating-point operations in the same
// order (par_cheby.c:318-371), with r, tmp never written for order 2 and u written once.
struct EpiChebyZero {
   static constexpr int NRED = 0;
   int scale; int last; int write_r; double ck; double ci; const double *f; const double *ds; double *rv;
   double *out; double *u;
   template <bool COH> __device__ __forceinline__ double xval(const double *__restrict__, int j) const
   {
      double res = ld_gather<COH>(f + j);
      if (scale) { res = MUL(ld_gather<COH>(ds + j), res); }
      const double u1 = MUL(res, ck);
      return scale ? MUL(ld_gather<COH>(ds + j), u1) : u1;
   }
   __device__ void operator()(int r, double t, double *) const
   {
      double res = f[r];
      if (scale) { res = MUL(ds[r], res); }
      if (write_r) { rv[r] = res; }
      double ui = scale ? ADD(MUL(ci, res), MUL(ds[r], t)) : ADD(MUL(ci, res), t);
      double su = scale ? MUL(ds[r], ui) : ui;
      if (last) { u[r] = su; }      // u == 0: 0 + su
      else      { out[r] = su; }
   }
};
// two-stage GS first stage (par_relax.c:1546-1557): z0 = (w*(f - t))/d
struct EpiGsFirst {
   static constexpr int NRED = 0;
   EPI_PLAIN_X
   double w; const double *f; const double *d; double *z;
   __device__ void operator()(int r, double t, double *) const { z[r] = __ddiv_rn(MUL(w, SUB(f[r], t)), d[r]); }
};
// inner stage over the strict lower triangle (par_relax.c:1562-1585): z' = t/d ; u += mult*z'
// (first inner stage also applies the pending u += z0)
template <bool DOT>
struct EpiGsInnerT {
   static constexpr int NRED = DOT ? 1 : 0;
   EPI_PLAIN_X
   int first; double mult; const double *d; const double *z_in; double *z_out; double *u; const double *dw;
   __device__ void operator()(int r, double t, double *red) const
   {
      double zn = __ddiv_rn(t, d[r]);
      double uu = u[r];
      if (first) { uu = ADD(uu, z_in[r]); }
      const double un = ADD(uu, MUL(mult, zn));
      u[r] = un;
      z_out[r] = zn;
      if (DOT) { red[0] = fma(dw[r], un, red[0]); }
   }
};

// zero-initial-guess / alpha == 0 elementwise variants (A*0 == 0: same arithmetic, no SpMV)
struct FScaleCopy { static constexpr int NRED = 0; __device__ void init() {} double beta; const double *b; double *y;
   __device__ void operator()(int i, double *) const { y[i] = MUL(beta, b[i]); } };
struct FZero { static constexpr int NRED = 0; __device__ void init() {} double w; const double *f; const double *l1; double *u;
   const int *marker; int points;
   __device__ void operator()(int i, double *) const
   {
      u[i] = (marker == nullptr || marker[i] == points) ? __ddiv_rn(MUL(w, f[i]), l1[i]) : 0.0;     // the others keep their zero
   } };
struct FChebyFirst0 { static constexpr int NRED = 0; __device__ void init() {} int scale; double c; const double *f; const double *ds;
   double *rv; double *tmp;
   __device__ void operator()(int i, double *) const
   {
      double res = f[i];
      if (scale) { res = MUL(ds[i], res); }
      rv[i] = res;
      double u1 = MUL(res, c);
      tmp[i] = scale ? MUL(ds[i], u1) : u1;
   } };
struct FGsFirst0 { static constexpr int NRED = 0; __device__ void init() {} double w; const double *f; const double *d; double *z;
   __device__ void operator()(int i, double *) const { z[i] = __ddiv_rn(MUL(w, f[i]), d[i]); } };

template <> struct EpiHeavy<EpiChebyZero>    { static constexpr bool value = true; };
template <> struct EpiHeavy<EpiMatvecDotDir> { static constexpr bool value = true; };

// ---------------------------------------------------------------- persistent cycle kernel (tail.cuh)
template <> struct TailKind<EpiMatvec>            { static constexpr int value = TK_MATVEC; };
template <> struct TailKind<EpiJacobiT<false>>    { static constexpr int value = TK_JACOBI; };
template <> struct TailKind<EpiChebyFirst>        { static constexpr int value = TK_CHEBY_FIRST; };
template <> struct TailKind<EpiChebyStepT<false>> { static constexpr int value = TK_CHEBY_STEP; };
template <> struct TailKind<EpiChebyZero>         { static constexpr int value = TK_CHEBY_ZERO; };
template <> struct TailKind<EpiGsFirst>           { static constexpr int value = TK_GS_FIRST; };
template <> struct TailKind<EpiGsInnerT<false>>   { static constexpr int value = TK_GS_INNER; };
template <> struct TailKind<FScaleCopy>           { static constexpr int value = TK_E_SCALECOPY; };
template <> struct TailKind<FZero>                { static constexpr int value = TK_E_ZERO; };
template <> struct TailKind<FChebyFirst0>         { static constexpr int value = TK_E_CHEBY_FIRST0; };
template <> struct TailKind<FGsFirst0>            { static constexpr int value = TK_E_GS_FIRST0; };

constexpr int TAIL_BLOCK = 256;
constexpr int TAIL_BLOCKS_PER_SM = 4;
constexpr long long TAIL_SPIN_LIMIT = 4000000000LL;     // ~2 s: a barrier that cannot complete reports instead of hanging

// rows of a CSR block with `lanes` lanes per row (a power of two, runtime), grid-strided over the whole grid.
// Loop order and shuffle tree are those of rowsum_csr<T>: bit-identical to the stand-alone launch.
template <class Epi>
__device__ __forceinline__ void tail_rows(const TailOp &op)
{
   Epi e;
   memcpy(&e, op.fn, sizeof(Epi));
   epi_init(e, 0);
   const int T = op.lanes;
   const int sub = threadIdx.x & (T - 1);
   const int rpb = TAIL_BLOCK / T;
   const int rpw = 32 / T;
   const int n_pad = (op.n + rpw - 1) / rpw * rpw;          // whole warps take part in the shuffles
   double red[1] = {0.0};
   for (int row = blockIdx.x * rpb + threadIdx.x / T; row < n_pad; row += gridDim.x * rpb)
   {
      const bool live = row < op.n;
      const int r = live ? row : op.n - 1;
      const int beg = op.rp[r], end = op.rp[r + 1];
      double t = 0.0;
      int k = beg + sub;
      for (; k + T < end; k += 2 * T)
      {
         const int c0 = ld_stream(op.ci + k), c1 = ld_stream(op.ci + k + T);
         const double v0 = ld_stream(op.va + k), v1 = ld_stream(op.va + k + T);
         const double x0 = e.template xval<true>(op.x, c0), x1 = e.template xval<true>(op.x, c1);
         t = fma(v0, x0, t); t = fma(v1, x1, t);
      }
      if (k < end)
      {
         const int c0 = ld_stream(op.ci + k);
         t = fma(ld_stream(op.va + k), e.template xval<true>(op.x, c0), t);
      }
      for (int o = T >> 1; o > 0; o >>= 1) { t += __shfl_down_sync(0xffffffffu, t, o, T); }
      if (live && sub == 0) { e(row, t, red); }
   }
}

template <class F>
__device__ __forceinline__ void tail_ewise(const TailOp &op)
{
   F f;
   memcpy(&f, op.fn, sizeof(F));
   f.init();
   double red[1] = {0.0};
   for (int i = blockIdx.x * TAIL_BLOCK + threadIdx.x; i < op.n; i += gridDim.x * TAIL_BLOCK) { f(i, red); }
}

// grid-wide barrier: arrival counter + generation word; the last block to arrive re-arms the counter and bumps the
// generation.  Every block reads the generation before its first arrival, so it cannot miss a bump.
__device__ __forceinline__ void tail_barrier(unsigned *bar, unsigned &gen)
{
   __syncthreads();
   if (threadIdx.x == 0)
   {
      __threadfence();
      if (atomicAdd(bar, 1u) == gridDim.x - 1)
      {
         bar[0] = 0u;
         __threadfence();
         atomicAdd(bar + 1, 1u);
      }
      else
      {
         const long long t0 = clock64();
         while (*(volatile unsigned *)(bar + 1) == gen)
         {
            if (clock64() - t0 > TAIL_SPIN_LIMIT) { atomicExch(bar + 2, 1u); break; }
         }
      }
      __threadfence();       // acquire side: also drops this SM's L1 lines (vectors rewritten by other SMs)
   }
   gen++;
   __syncthreads();
}

__global__ void __launch_bounds__(TAIL_BLOCK, TAIL_BLOCKS_PER_SM)
tail_kernel(const TailOp *__restrict__ ops, int n_ops, unsigned *bar, const int *stop)
{
   if (stop != nullptr && *stop != 0) { return; }
   __shared__ TailOp s_op;
   unsigned gen = *(volatile unsigned *)(bar + 1);
   for (int k = 0; k < n_ops; k++)
   {
      // (the previous barrier's trailing __syncthreads keeps s_op stable until every thread is done with it)
      for (int w = threadIdx.x; w < (int)(sizeof(TailOp) / 4); w += TAIL_BLOCK)
      {
         reinterpret_cast<unsigned *>(&s_op)[w] = reinterpret_cast<const unsigned *>(ops + k)[w];
      }
      __syncthreads();
      const TailOp &op = s_op;
      switch (op.kind)
      {
         case TK_MATVEC:         tail_rows<EpiMatvec>(op); break;
         case TK_JACOBI:         tail_rows<EpiJacobiT<false>>(op); break;
         case TK_CHEBY_FIRST:    tail_rows<EpiChebyFirst>(op); break;
         case TK_CHEBY_STEP:     tail_rows<EpiChebyStepT<false>>(op); break;
         case TK_CHEBY_ZERO:     tail_rows<EpiChebyZero>(op); break;
         case TK_GS_FIRST:       tail_rows<EpiGsFirst>(op); break;
         case TK_GS_INNER:       tail_rows<EpiGsInnerT<false>>(op); break;
         case TK_E_SCALECOPY:    tail_ewise<FScaleCopy>(op); break;
         case TK_E_ZERO:         tail_ewise<FZero>(op); break;
         case TK_E_CHEBY_FIRST0: tail_ewise<FChebyFirst0>(op); break;
         case TK_E_GS_FIRST0:    tail_ewise<FGsFirst0>(op); break;
         case TK_FILL:  for (int i = blockIdx.x * TAIL_BLOCK + threadIdx.x; i < op.n; i += gridDim.x * TAIL_BLOCK) { op.y[i] = op.a; } break;
         case TK_COPY:  for (int i = blockIdx.x * TAIL_BLOCK + threadIdx.x; i < op.n; i += gridDim.x * TAIL_BLOCK) { op.y[i] = __ldcg(op.x + i); } break;
         case TK_AXPY:  for (int i = blockIdx.x * TAIL_BLOCK + threadIdx.x; i < op.n; i += gridDim.x * TAIL_BLOCK) { op.y[i] = fma(op.a, __ldcg(op.x + i), op.y[i]); } break;
         case TK_SCALE: for (int i = blockIdx.x * TAIL_BLOCK + threadIdx.x; i < op.n; i += gridDim.x * TAIL_BLOCK) { op.y[i] *= op.a; } break;
         case TK_DENSE:
         {
            // u[i] = sum_j M[i][j] f[j]; one warp per row (dense_matvec_kernel's loop and shuffle tree)
            const int lane = threadIdx.x & 31;
            for (int row = blockIdx.x * (TAIL_BLOCK >> 5) + (threadIdx.x >> 5); row < op.n; row += gridDim.x * (TAIL_BLOCK >> 5))
            {
               double t = 0.0;
               for (int j = lane; j < op.n_cols; j += 32) { t = fma(op.va[(size_t)row * op.n_cols + j], __ldcg(op.x + j), t); }
#pragma unroll
               for (int o = 16; o > 0; o >>= 1) { t += __shfl_down_sync(0xffffffffu, t, o); }
               if (lane == 0) { op.y[row] = t; }
            }
            break;
         }
         default: break;
      }
      if (k + 1 < n_ops) { tail_barrier(bar, gen); }
   }
}

struct TailProgram { std::vector<TailOp> ops; TailOp *d_ops = nullptr; };
struct TailRec {
   bool recording = false;
   std::vector<TailOp> cur;
   std::vector<TailProgram> cache;
   unsigned *d_bar = nullptr;         // [0] arrivals, [1] generation, [2] barrier timed out
   int grid = 0;
   int coop = 1;
};

static TailRec *tail_of(b200_ctx *c)
{
   if (!c->tail) { c->tail = new TailRec(); }
   return static_cast<TailRec *>(c->tail);
}

bool tail_recording(const b200_ctx *c) { return c->tail != nullptr && static_cast<const TailRec *>(c->tail)->recording; }

int tail_begin(b200_ctx *c)
{
   TailRec *t = tail_of(c);
   t->cur.clear();
   t->recording = true;
   return B200_OK;
}

int tail_append(b200_ctx *c, const TailOp &op)
{
   tail_of(c)->cur.push_back(op);
   return B200_OK;
}

void tail_drop_programs(b200_ctx *c)
{
   if (!c->tail) { return; }
   TailRec *t = static_cast<TailRec *>(c->tail);
   for (auto &p : t->cache) { cudaFree(p.d_ops); }
   t->cache.clear();
   t->recording = false;
   t->cur.clear();
}

void tail_release(b200_ctx *c)
{
   if (!c->tail) { return; }
   tail_drop_programs(c);
   TailRec *t = static_cast<TailRec *>(c->tail);
   cudaFree(t->d_bar);
   delete t;
   c->tail = nullptr;
}

int tail_flush(b200_ctx *c)
{
   TailRec *t = tail_of(c);
   t->recording = false;
   if (t->cur.empty()) { return B200_OK; }
   if (t->d_bar == nullptr)
   {
      CUDA_TRY(cudaMalloc((void **)&t->d_bar, 4 * sizeof(unsigned)));
      CUDA_TRY(cudaMemset(t->d_bar, 0, 4 * sizeof(unsigned)));
      int per_sm = 0;
      CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tail_kernel, TAIL_BLOCK, 0));
      if (per_sm < 1) { b200_set_error("persistent cycle kernel does not fit an SM"); return B200_ERR_CUDA; }
      if (per_sm > TAIL_BLOCKS_PER_SM) { per_sm = TAIL_BLOCKS_PER_SM; }
      int sms = B200_NUM_SMS;
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
      t->grid = sms * per_sm;
      const char *e = getenv("B200_TAIL_COOP");
      t->coop = e ? atoi(e) : 1;
   }
   // the program of a cycle is the same every time: upload once, find it again by content
   TailProgram *prog = nullptr;
   for (auto &p : t->cache)
   {
      if (p.ops.size() == t->cur.size() && memcmp(p.ops.data(), t->cur.data(), sizeof(TailOp) * t->cur.size()) == 0) { prog = &p; break; }
   }
   if (prog == nullptr)
   {
      if (c->capturing)
      {
         b200_set_error("persistent cycle kernel: new program during graph capture");
         return B200_ERR_CUDA;
      }
      if (t->cache.size() >= 64) { tail_drop_programs(c); }
      TailProgram p;
      p.ops = t->cur;
      CUDA_TRY(cudaMalloc((void **)&p.d_ops, sizeof(TailOp) * p.ops.size()));
      CUDA_TRY(cudaMemcpyAsync(p.d_ops, p.ops.data(), sizeof(TailOp) * p.ops.size(), cudaMemcpyHostToDevice, c->stream));
      CUDA_TRY(cudaStreamSynchronize(c->stream));      // (first cycle only; the host vector is about to move)
      t->cache.push_back(std::move(p));
      prog = &t->cache.back();
   }
   const TailOp *d_ops = prog->d_ops;
   int n_ops = (int)prog->ops.size();
   unsigned *bar = t->d_bar;
   const int *stop = c->d_stop;
   if (c->prof) { prof_begin(c, "tail_kernel (persistent coarse-level cycle)"); }
   if (t->coop)
   {
      void *args[] = {(void *)&d_ops, (void *)&n_ops, (void *)&bar, (void *)&stop};
      CUDA_TRY(cudaLaunchCooperativeKernel((const void *)tail_kernel, dim3(t->grid), dim3(TAIL_BLOCK), args, 0, c->stream));
   }
   else
   {
      tail_kernel<<<t->grid, TAIL_BLOCK, 0, c->stream>>>(d_ops, n_ops, bar, stop);
   }
   if (c->prof) { prof_end(c); }
   c->launches++;
   CHECK_LAUNCH();
   t->cur.clear();
   return B200_OK;
}

// 0 = fine, 1 = a grid barrier of the persistent cycle kernel timed out (results invalid)
int tail_error(b200_ctx *c)
{
   if (!c->tail) { return 0; }
   TailRec *t = static_cast<TailRec *>(c->tail);
   if (!t->d_bar) { return 0; }
   unsigned e = 0;
   if (cudaMemcpy(&e, t->d_bar + 2, sizeof(unsigned), cudaMemcpyDeviceToHost) != cudaSuccess) { return 1; }
   return e != 0;
}

// ---------------------------------------------------------------- operators
int op_matvec(b200_parcsr *A, double alpha, const double *x, double beta, const double *b, double *y)
{
   if (alpha == 0.0)
   {
      // y = beta*b (csr_matvec.c:130-160)
      if (beta == 0.0) { return vec_fill(A->ctx, A->n_rows, 0.0, y); }
      FScaleCopy f{beta, b, y};
      return run_ewise(A->ctx, A->n_rows, f, nullptr);
   }
   EpiMatvec e{alpha, beta, b, y};
   return run_rows(A, A->diag, A->offd, true, x, e, nullptr);
}

int op_matvec_dot(b200_parcsr *A, const double *x, double *y, double *d_out)
{
   EpiMatvecDot e{x, y};
   return run_rows(A, A->diag, A->offd, true, x, e, d_out);
}

bool op_can_fuse_gather(const b200_parcsr *A)
{
   // (the NCCL baseline path packs the raw vector with its own kernel: no composite operands there)
   return (A->num_sends + A->num_recvs) == 0 || A->ch_fwd != nullptr;
}

int op_matvec_dot_direction(b200_parcsr *A, const double *d_beta, const int *d_iter, const double *z,
                            double *pbuf0, double *pbuf1, double *s, double *d_out)
{
   EpiMatvecDotDir e{d_beta, d_iter, z, pbuf0, pbuf1, s, 0.0, nullptr, nullptr};
   return run_rows(A, A->diag, A->offd, true, z, e, d_out);
}

struct FScaleInPlace { static constexpr int NRED = 0; __device__ void init() {} double beta; double *y;
   __device__ void operator()(int i, double *) const { y[i] = beta * y[i]; } };

// hypre_ParCSRMatrixMatvecTHost (src/parcsr_mv/par_csr_matvec.c:287-514):
//   y_tmp = offd^T x -> owners (job 2)  ||  y = alpha diag^T x + beta y  ; y[send_map] += y_buf
int op_matvecT(b200_parcsr *A, double alpha, const double *x, double beta, double *y)
{
   B200_TRY(parcsr_ensure_T(A));
   b200_ctx *c = A->ctx;
   const bool comm = c->nranks > 1 && (A->num_sends + A->num_recvs) > 0;
   if (comm && A->ch_rev)
   {
      // peer-memory path: offd^T x -> owners, diag^T x and the deterministic add in one launch
      EpiMatvec e{alpha, beta, y, y};
      return run_rows_transposed_p2p(A, x, e);
   }
   if (comm && A->n_cols_offd > 0)
   {
      EpiMatvec eo{1.0, 0.0, nullptr, A->d_ghost_out};
      B200_TRY(run_rows(A, A->offdT, A->offdT, false, x, eo, nullptr));
   }
   if (comm) { B200_TRY(halo_reverse_start(A)); }
   EpiMatvec e{alpha, beta, y, y};
   B200_TRY(run_rows(A, A->diagT, A->offdT, false, x, e, nullptr));
   if (comm) { B200_TRY(halo_reverse_finish_add(A, alpha, y)); }
   return B200_OK;
}

int op_l1jacobi(b200_parcsr *A, const double *f, const double *l1, double w,
                const double *u_in, double *u_out, int zero_guess, const FusedDot *fd, const int *marker, int points)
{
   if (points == 0) { marker = nullptr; }
   if (zero_guess)
   {
      // Vtemp = w*f ; u = 0 + Vtemp./l1 (par_relax.c:1221-1228,1238-1246)
      FZero fz{w, f, l1, u_out, marker, points};
      return run_ewise(A->ctx, A->n_rows, fz, nullptr);
   }
   if (fd && fd->with)
   {
      EpiJacobiT<true> e{w, f, l1, u_in, u_out, fd->with, marker, points};
      *fd->done = true;
      return run_rows(A, A->diag, A->offd, true, u_in, e, fd->out);
   }
   EpiJacobiT<false> e{w, f, l1, u_in, u_out, nullptr, marker, points};
   return run_rows(A, A->diag, A->offd, true, u_in, e, nullptr);
}

int op_cheby(b200_parcsr *A, const double *f, const double *ds, const double *coefs_h, int order,
             int scale, double *u, double *v, double *r, double *tmp, int zero_guess, const FusedDot *fd)
{
   if (order > 4) { order = 4; }
   if (order < 1) { order = 1; }
   const int k = order - 1;   // cheby_order (par_cheby.c:262)
   b200_ctx *c = A->ctx;
   const int n = A->n_rows;
   int i_first = k - 1;
   double *in = tmp, *out = v;
   const bool can_fuse = op_can_fuse_gather(A);
   if (zero_guess && k >= 1 && can_fuse)
   {
      // first elementwise step fused into the first polynomial SpMV (see EpiChebyZero)
      const int last = (k == 1) ? 1 : 0;
      EpiChebyZero e{scale, last, k > 1 ? 1 : 0, coefs_h[k], coefs_h[k - 1], f, ds, r, out, u};
      B200_TRY(run_rows(A, A->diag, A->offd, true, f, e, nullptr));
      if (last) { return B200_OK; }
      double *t = in; in = out; out = t;
      i_first = k - 2;
   }
   else if (zero_guess)
   {
      FChebyFirst0 f0{scale, coefs_h[k], f, ds, r, tmp};
      B200_TRY(run_ewise(c, n, f0, nullptr));
   }
   else
   {
      EpiChebyFirst e{scale, coefs_h[k], f, ds, r, tmp};
      B200_TRY(run_rows(A, A->diag, A->offd, true, u, e, nullptr));
   }
   if (k == 0)
   {
      // u = orig_u + ds.*u1  (tmp already holds ds.*u1)
      return zero_guess ? b200_vec_copy(c, n, tmp, u) : vec_axpy(c, n, 1.0, tmp, u);
   }
   for (int i = i_first; i >= 0; i--)
   {
      if (i == 0 && zero_guess)
      {
         // u == 0: write instead of accumulate (same value: 0 + x == x)
         B200_TRY(vec_fill(c, n, 0.0, u));
      }
      if (i == 0 && fd && fd->with)
      {
         EpiChebyStepT<true> e{scale, 1, coefs_h[i], ds, r, out, u, fd->with};
         *fd->done = true;
         B200_TRY(run_rows(A, A->diag, A->offd, true, in, e, fd->out));
      }
      else
      {
         EpiChebyStepT<false> e{scale, i == 0 ? 1 : 0, coefs_h[i], ds, r, out, u, nullptr};
         B200_TRY(run_rows(A, A->diag, A->offd, true, in, e, nullptr));
      }
      double *t = in; in = out; out = t;
   }
   return B200_OK;
}

int op_gs2(b200_parcsr *A, const double *f, double w, int inner, double *u, double *z, double *tmp,
           int zero_guess, const FusedDot *fd)
{
   B200_TRY(parcsr_ensure_L(A));
   if (A->zero_diagonal)
   {
      b200_set_error("two-stage Gauss-Seidel: zero diagonal entry (reference: hypre_error_in_arg(1), par_relax.c:1538-1544)");
      return B200_ERR_NUMERIC;
   }
   b200_ctx *c = A->ctx;
   const int n = A->n_rows;
   const double *d = A->d_diagonal;
   if (zero_guess)
   {
      FGsFirst0 f0{w, f, d, z};
      B200_TRY(run_ewise(c, n, f0, nullptr));
      B200_TRY(vec_fill(c, n, 0.0, u));
   }
   else
   {
      EpiGsFirst e{w, f, d, z};
      B200_TRY(run_rows(A, A->diag, A->offd, true, u, e, nullptr));
   }
   if (inner <= 0) { return vec_axpy(c, n, 1.0, z, u); }
   double mult = -1.0;
   double *zin = z, *zout = tmp;
   for (int k = 0; k < inner; k++)
   {
      if (k == inner - 1 && fd && fd->with)
      {
         EpiGsInnerT<true> e{k == 0 ? 1 : 0, mult, d, zin, zout, u, fd->with};
         *fd->done = true;
         B200_TRY(run_rows(A, A->lower, A->lower, false, zin, e, fd->out));
      }
      else
      {
         EpiGsInnerT<false> e{k == 0 ? 1 : 0, mult, d, zin, zout, u, nullptr};
         B200_TRY(run_rows(A, A->lower, A->lower, false, zin, e, nullptr));
      }
      mult = -mult;
      double *t = zin; zin = zout; zout = t;
   }
   return B200_OK;
}

// ---------------------------------------------------------------- dense coarse solve
// u[i] = sum_j Ainv[i][j] * f_all[j]; one warp per row.
__global__ void dense_matvec_kernel(int n_rows, int n_cols, const double *__restrict__ M,
                                    const double *__restrict__ f, double *__restrict__ u)
{
   const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
   const int lane = threadIdx.x & 31;
   if (row >= n_rows) { return; }
   double t = 0.0;
   for (int j = lane; j < n_cols; j += 32) { t = fma(M[(size_t)row * n_cols + j], f[j], t); }
#pragma unroll
   for (int o = 16; o > 0; o >>= 1) { t += __shfl_down_sync(0xffffffffu, t, o); }
   if (lane == 0) { u[row] = t; }
}

int op_dense_matvec(b200_ctx *ctx, int n_rows, int n_cols, const double *M, const double *f_all, double *u)
{
   if (n_rows == 0) { return B200_OK; }
   if (tail_recording(ctx))
   {
      TailOp op;
      memset(&op, 0, sizeof(op));
      op.kind = TK_DENSE; op.n = n_rows; op.n_cols = n_cols; op.va = M; op.x = f_all; op.y = u;
      return tail_append(ctx, op);
   }
   int grid = (n_rows + 7) / 8;
   LAUNCH(ctx, dense_matvec_kernel, grid, 256, 0, n_rows, n_cols, M, f_all, u);
   CHECK_LAUNCH();
   return B200_OK;
}

// ---------------------------------------------------------------- C ABI
extern "C" int b200_parcsr_matvec(b200_parcsr *A, double alpha, const double *x, double beta,
                                  const double *b, double *y)
{
   ARG_CHECK(A != nullptr && y != nullptr && (x != nullptr || A->n_cols_diag == 0));
   ARG_CHECK(beta == 0.0 || b != nullptr);
   ARG_CHECK(x != y);
   return op_matvec(A, alpha, x, beta, b, y);
}

extern "C" int b200_parcsr_matvecT(b200_parcsr *A, double alpha, const double *x, double beta, double *y)
{
   ARG_CHECK(A != nullptr && y != nullptr && x != y);
   return op_matvecT(A, alpha, x, beta, y);
}

extern "C" int b200_relax_l1jacobi(b200_parcsr *A, const double *f, const double *l1, double w,
                                   double *u, double *tmp, int zero_guess)
{
   ARG_CHECK(A != nullptr && f != nullptr && l1 != nullptr && u != nullptr && tmp != nullptr);
   if (zero_guess) { return op_l1jacobi(A, f, l1, w, u, u, 1, nullptr, nullptr, 0); }
   B200_TRY(op_l1jacobi(A, f, l1, w, u, tmp, 0, nullptr, nullptr, 0));
   return b200_vec_copy(A->ctx, A->n_rows, tmp, u);
}

extern "C" int b200_relax_cheby(b200_parcsr *A, const double *f, const double *ds, const double *coefs_h,
                                int order, int scale, double *u, double *v, double *r, double *tmp,
                                int zero_guess)
{
   ARG_CHECK(A != nullptr && f != nullptr && coefs_h != nullptr && u != nullptr && v != nullptr &&
             r != nullptr && tmp != nullptr);
   ARG_CHECK(!scale || ds != nullptr);
   return op_cheby(A, f, ds, coefs_h, order, scale, u, v, r, tmp, zero_guess, nullptr);
}

extern "C" int b200_relax_2stage_gs(b200_parcsr *A, const double *f, double w, int inner,
                                    double *u, double *z, double *tmp, int zero_guess)
{
   ARG_CHECK(A != nullptr && f != nullptr && u != nullptr && z != nullptr && tmp != nullptr && inner >= 1);
   return op_gs2(A, f, w, inner, u, z, tmp, zero_guess, nullptr);
}

// host-vector SpMV: what HYPRE_ParCSRMatrixMatvec costs end to end for host-resident
// hypre vectors (H2D x [, b], kernel, D2H y).
extern "C" int b200_parcsr_matvec_host(b200_parcsr *A, double alpha, const double *x_h, double beta,
                                       const double *b_h, double *y_h)
{
   ARG_CHECK(A != nullptr && x_h != nullptr && y_h != nullptr && (beta == 0.0 || b_h != nullptr));
   b200_ctx *c = A->ctx;
   double *dx = nullptr, *dy = nullptr;
   B200_TRY(ctx_workspace(c, WS_HOST_X, A->n_cols_diag, &dx));
   B200_TRY(ctx_workspace(c, WS_HOST_Y, A->n_rows, &dy));
   CUDA_TRY(cudaMemcpyAsync(dx, x_h, sizeof(double) * (size_t)A->n_cols_diag, cudaMemcpyHostToDevice, c->stream));
   if (beta != 0.0)
   {
      CUDA_TRY(cudaMemcpyAsync(dy, b_h, sizeof(double) * (size_t)A->n_rows, cudaMemcpyHostToDevice, c->stream));
   }
   B200_TRY(op_matvec(A, alpha, dx, beta, dy, dy));
   CUDA_TRY(cudaMemcpyAsync(y_h, dy, sizeof(double) * (size_t)A->n_rows, cudaMemcpyDeviceToHost, c->stream));
   CUDA_TRY(cudaStreamSynchronize(c->stream));
   return B200_OK;
}

// host-vector SpMV^T: what HYPRE_ParCSRMatrixMatvecT costs end to end for host-resident hypre vectors
// (H2D x [, y], kernels, D2H y).  x has n_rows entries, y this rank's n_cols_diag.
extern "C" int b200_parcsr_matvecT_host(b200_parcsr *A, double alpha, const double *x_h, double beta, double *y_h)
{
   ARG_CHECK(A != nullptr && x_h != nullptr && y_h != nullptr);
   b200_ctx *c = A->ctx;
   double *dx = nullptr, *dy = nullptr;
   B200_TRY(ctx_workspace(c, WS_HOST_X, A->n_rows, &dx));
   B200_TRY(ctx_workspace(c, WS_HOST_Y, A->n_cols_diag, &dy));
   CUDA_TRY(cudaMemcpyAsync(dx, x_h, sizeof(double) * (size_t)A->n_rows, cudaMemcpyHostToDevice, c->stream));
   if (beta != 0.0)
   {
      CUDA_TRY(cudaMemcpyAsync(dy, y_h, sizeof(double) * (size_t)A->n_cols_diag, cudaMemcpyHostToDevice, c->stream));
   }
   B200_TRY(op_matvecT(A, alpha, dx, beta, dy));
   CUDA_TRY(cudaMemcpyAsync(y_h, dy, sizeof(double) * (size_t)A->n_cols_diag, cudaMemcpyDeviceToHost, c->stream));
   CUDA_TRY(cudaStreamSynchronize(c->stream));
   return B200_OK;
}
