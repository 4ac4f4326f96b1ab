// rowlaunch.cuh -- host-side dispatch of the fused row kernels.
#pragma once
#include "spmv.cuh"
#include "tail.cuh"
#include <cstring>
#include <algorithm>

static inline RedCtx make_redctx(b200_ctx *ctx, double *out, int accumulate)
{
   RedCtx rc;
   rc.partials = ctx->d_partials; rc.counter = ctx->d_counter; rc.out = out; rc.accumulate = accumulate;
   rc.stop = ctx->d_stop;
   return rc;
}

// grid for `work` row-groups of which `per_block` fit a block: whole waves of the
// 148 SMs x 8 resident 256-thread blocks, folded by the kernels' stride loops
extern int g_rows_waves;
static inline int rows_grid(long long work, int per_block, int waves = 4)
{
   long long g = (work + per_block - 1) / per_block;
   const long long cap = (long long)B200_NUM_SMS * 8 * waves;   // 4736 <= B200_MAX_PARTIALS
   if (g > cap) { g = cap; }
   if (g < 1) { g = 1; }
   return (int)g;
}

// launch options of the thread-per-row kernels (pack.cu; environment B200_CHUNKED, B200_PAT_REGS,
// b200_set_kernel_options): slice-to-block mapping and the PAT8 instantiation
extern int g_rows_chunked, g_pat_regs;

static inline HaloArgs no_halo()
{
   HaloArgs h;
   memset(&h, 0, sizeof(h));
   return h;
}

template <int HALO, class Epi>
static int launch_rows_layout(b200_ctx *ctx, const DMat &D, const HaloArgs &h, const double *x,
                              const Epi &epi, const RedCtx &rc)
{
   MatView V = view_of(D);
   constexpr bool P2P = (HALO == HALO_FWD || HALO == HALO_REV);
   if (tail_recording(ctx))
   {
      // persistent cycle kernel: record instead of launching (CSR-vector blocks without communication only)
      if (HALO != HALO_NONE || D.layout != 0)
      {
         b200_set_error("operator with communication or a thread-per-row layout inside the persistent cycle kernel");
         return B200_ERR_UNSUPPORTED;
      }
      if (D.n_rows == 0) { return B200_OK; }
      TailOp op;
      memset(&op, 0, sizeof(op));
      op.n = D.n_rows; op.lanes = D.lanes; op.rp = D.rp; op.ci = D.ci; op.va = D.va; op.x = x;
      return tail_record_fn(ctx, op, epi);
   }
   if (D.n_rows == 0 && !P2P)
   {
      // reductions still have to publish a zero
      if (Epi::NRED > 0 && !rc.accumulate) { CUDA_TRY(cudaMemsetAsync(rc.out, 0, sizeof(double) * Epi::NRED, ctx->stream)); }
      return B200_OK;
   }
   // (a rank without rows still launches in the peer-memory modes: its blocks carry the put phase
   //  and the reduction publishes the zero)
   if (D.layout >= 1 || D.n_rows == 0)
   {
      int grid = rows_grid(D.n_slices, ROWK_BLOCK / 32, D.cfmt == CFMT_PAT8 ? g_rows_waves : 4);
      HaloArgs hh = h;
      hh.chunked = D.cfmt == CFMT_PAT8 ? g_rows_chunked : (g_rows_chunked > 1);
      const int variant = D.cfmt == CFMT_PAT8 ? 0 : (D.zpad ? 1 : 0);
      switch (D.cfmt ? D.cfmt + 8 * variant : 0)
      {
#define CASE_C(TT) case TT: LAUNCH(ctx, (rows_kernel<-(TT), HALO, Epi>), grid, ROWK_BLOCK, 0, V, hh, x, epi, rc); break;
         CASE_C(CFMT_VAL8) CASE_C(CFMT_VAL16) CASE_C(CFMT_VAL64) CASE_C(CFMT_PAT8)
         CASE_C(CFMT_VAL8 + 8) CASE_C(CFMT_VAL16 + 8) CASE_C(CFMT_VAL64 + 8)
#undef CASE_C
         default: LAUNCH(ctx, (rows_kernel<0, HALO, Epi>), grid, ROWK_BLOCK, 0, V, hh, x, epi, rc); break;
      }
   }
   else
   {
      switch (D.lanes)
      {
#define CASE_T(TT)                                                                            \
   case TT: {                                                                                 \
      int grid = rows_grid(D.n_rows, ROWK_BLOCK / TT);                                        \
      LAUNCH(ctx, (rows_kernel<TT, HALO, Epi>), grid, ROWK_BLOCK, 0, V, h, x, epi, rc);       \
   } break;
         CASE_T(1) CASE_T(2) CASE_T(4) CASE_T(8) CASE_T(16) CASE_T(32)
#undef CASE_T
         default:
            b200_set_error("unsupported lanes_per_row %d", D.lanes);
            return B200_ERR_ARG;
      }
   }
   CHECK_LAUNCH();
   return B200_OK;
}

// t_i = (diag x_local + offd x_ghost)_i for every local row i, then epi(i, t_i).
// Multi-rank, peer-memory path: ONE launch (rows_kernel<.., HALO_FWD, ..>: pack + put, rows,
// lazy wait, acknowledge).  Multi-rank, NCCL path: pack+exchange on the comm stream, interior rows
// meanwhile, boundary rows after the ghosts landed
// (hypre_ParCSRMatrixMatvecOutOfPlaceHost, src/parcsr_mv/par_csr_matvec.c:163-218).
// red_out: device scalars receiving the Epi::NRED fused reductions (local sums).
template <class Epi>
static int run_rows(b200_parcsr *A, const DMat &D, const DMat &O, bool use_halo,
                    const double *x, const Epi &epi, double *red_out)
{
   b200_ctx *ctx = A->ctx;
   if (use_halo && (A->num_sends + A->num_recvs) > 0)
   {
      if (tail_recording(ctx))
      {
         b200_set_error("operator with a halo exchange inside the persistent cycle kernel");
         return B200_ERR_UNSUPPORTED;
      }
      HaloArgs h = no_halo();
      h.O = view_of(O);
      if (A->ch_fwd)
      {
         h.ch = p2p_chan_view(A->ch_fwd);
         h.map = A->d_send_map;
         h.grp = A->d_grp_fwd;
         h.int_begin = A->int_fwd[0]; h.int_end = A->int_fwd[1];
         h.dbg = ctx->p2p_debug;
         return launch_rows_layout<HALO_FWD, Epi>(ctx, D, h, x, epi, make_redctx(ctx, red_out, 0));
      }
      B200_TRY(halo_start(A, x));
      B200_TRY((launch_rows_layout<HALO_SKIP, Epi>(ctx, D, h, x, epi, make_redctx(ctx, red_out, 0))));
      B200_TRY(halo_finish(A));
      if (A->n_bnd > 0)
      {
         int grid = rows_grid(A->n_bnd, ROWK_BLOCK / 4);
         LAUNCH(ctx, (rowlist_kernel<Epi>), grid, ROWK_BLOCK, 0, view_of(D), view_of(O),
                A->d_bnd_rows, A->n_bnd, x, A->d_ghost, epi, make_redctx(ctx, red_out, 1));
         CHECK_LAUNCH();
      }
      return B200_OK;
   }
   return launch_rows_layout<HALO_NONE, Epi>(ctx, D, no_halo(), x, epi, make_redctx(ctx, red_out, 0));
}

// y = alpha*diag^T x + beta*y + alpha*(contributions of the other ranks), peer-memory path, one launch
template <class Epi>
static int run_rows_transposed_p2p(b200_parcsr *A, const double *x, const Epi &epi)
{
   HaloArgs h = no_halo();
   h.O = view_of(A->offdT);
   h.ch = p2p_chan_view(A->ch_rev);
   h.urp = A->d_unpack_rp_dense;
   h.upos = A->d_unpack_pos;
   h.grp = A->d_grp_rev;
   h.int_begin = A->int_rev[0]; h.int_end = A->int_rev[1];
   h.dbg = A->ctx->p2p_debug;
   return launch_rows_layout<HALO_REV, Epi>(A->ctx, A->diagT, h, x, epi, make_redctx(A->ctx, nullptr, 0));
}
