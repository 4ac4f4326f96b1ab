// tail.cuh -- the coarse part of a multigrid cycle as ONE persistent kernel.
//
// On the coarse levels of a hierarchy (hypre_BoomerAMGCycle, src/parcsr_ls/par_cycle.c:209-779, levels with a
// few thousand rows) every operator is a launch of a few microseconds of which most is launch latency: the
// 256^3 Chebyshev V-cycle spends ~50 of its ~75 launches there.  Between tail_begin() and tail_flush() the
// launchers (rowlaunch.cuh, vecops.cuh, vecops.cu, ops.cu) do not launch: they append a TailOp -- the
// operator's matrix, vectors and its epilogue functor, byte for byte -- to the context's recorder; tail_flush()
// uploads the program (once: it is identical from one cycle to the next) and launches tail_kernel, a cooperative
// grid of 148 x 4 blocks that executes the operators in order with a grid-wide barrier between them.  Same
// epilogues, same lane counts, same summation order as the stand-alone launches: results are bit-identical.
// Only operators without communication and without reductions qualify (coarse levels of a single-rank solve,
// replicated coarse levels of a multi-rank solve).
#pragma once
#include "common.cuh"

enum TailOpKind {
   TK_NONE = -1,
   // row operators (epilogue functor in TailOp::fn), defined in ops.cu
   TK_MATVEC = 0, TK_JACOBI, TK_CHEBY_FIRST, TK_CHEBY_STEP, TK_CHEBY_ZERO, TK_GS_FIRST, TK_GS_INNER,
   // elementwise operators (functor in TailOp::fn), defined in ops.cu
   TK_E_SCALECOPY, TK_E_ZERO, TK_E_CHEBY_FIRST0, TK_E_GS_FIRST0,
   // BLAS-1 recorded by value: y = a (FILL), y = x (COPY), y += a x (AXPY), y *= a (SCALE)
   TK_FILL, TK_COPY, TK_AXPY, TK_SCALE,
   // dense coarse solve u = M f (rows x n_cols)
   TK_DENSE,
};

constexpr int TAIL_FN_BYTES = 104;

struct TailOp {
   int kind, n, lanes, n_cols;
   const int *rp, *ci;
   const double *va;            // CSR values | dense matrix
   const double *x;             // gathered operand | x of COPY / AXPY | f of DENSE
   double *y;                   // FILL / COPY / AXPY / SCALE / DENSE result
   double a;
   alignas(8) unsigned char fn[TAIL_FN_BYTES];
};

// which TailOpKind a functor type is recorded as (specialised next to the functors, ops.cu); TK_NONE = the
// operator cannot be part of a recorded program
template <class F> struct TailKind { static constexpr int value = TK_NONE; };

// recorder API (ops.cu)
bool tail_recording(const b200_ctx *c);
int  tail_begin(b200_ctx *c);
int  tail_flush(b200_ctx *c);                       // launch what was recorded since tail_begin, stop recording
void tail_release(b200_ctx *c);                     // free the recorder (context destruction)
void tail_drop_programs(b200_ctx *c);               // forget cached programs (their vectors / matrices are going away)
int  tail_append(b200_ctx *c, const TailOp &op);
int  tail_error(b200_ctx *c);                       // a barrier timed out (checked after a solve)

template <class F>
static inline int tail_record_fn(b200_ctx *c, TailOp op, const F &f)
{
   static_assert(sizeof(F) <= TAIL_FN_BYTES, "functor too large for a TailOp");
   static_assert(std::is_trivially_copyable<F>::value, "functor must be trivially copyable");
   if (TailKind<F>::value == TK_NONE || F::NRED != 0)
   {
      b200_set_error("operator cannot run inside the persistent cycle kernel (functor not registered)");
      return B200_ERR_UNSUPPORTED;
   }
   op.kind = TailKind<F>::value;
   memset(op.fn, 0, sizeof(op.fn));
   memcpy(op.fn, &f, sizeof(F));
   return tail_append(c, op);
}
