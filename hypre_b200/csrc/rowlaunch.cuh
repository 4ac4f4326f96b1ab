// rowlaunch.cuh -- host-side dispatch of the fused row kernels.
#pragma once
#include "spmv.cuh"

static inline RedCtx make_redctx(b200_ctx *ctx, double *out, int accumulate)
{
   RedCtx rc;
   rc.partials = ctx->d_partials; rc.counter = ctx->d_counter; rc.out = out; rc.accumulate = accumulate;
   return rc;
}

// grid for `work` row-groups of which `per_block` fit a block: whole waves of the
// 148 SMs x 8 resident 256-thread blocks, folded by the kernels' stride loops
static inline int rows_grid(long long work, int per_block)
{
   long long g = (work + per_block - 1) / per_block;
   const long long cap = (long long)B200_NUM_SMS * 8 * 4;   // 4736 <= B200_MAX_PARTIALS
   if (g > cap) { g = cap; }
   if (g < 1) { g = 1; }
   return (int)g;
}

template <bool SKIP, class Epi>
static int launch_rows_layout(b200_ctx *ctx, const DMat &D, const int *offd_rp, const double *x,
                              const Epi &epi, const RedCtx &rc)
{
   MatView V = view_of(D);
   if (D.n_rows == 0)
   {
      // reductions still have to publish a zero
      if (Epi::NRED > 0 && !rc.accumulate) { CUDA_TRY(cudaMemsetAsync(rc.out, 0, sizeof(double) * Epi::NRED, ctx->stream)); }
      return B200_OK;
   }
   if (D.layout == 1)
   {
      int grid = rows_grid(D.n_slices, ROWK_BLOCK / 32);
      LAUNCH(ctx, (rows_kernel<0, SKIP, Epi>), grid, ROWK_BLOCK, 0, V, offd_rp, x, epi, rc);
   }
   else
   {
      switch (D.lanes)
      {
#define CASE_T(TT)                                                                            \
   case TT: {                                                                                 \
      int grid = rows_grid(D.n_rows, ROWK_BLOCK / TT);                                        \
      LAUNCH(ctx, (rows_kernel<TT, SKIP, Epi>), grid, ROWK_BLOCK, 0, V, offd_rp, x, epi, rc); \
   } break;
         CASE_T(2) CASE_T(4) CASE_T(8) CASE_T(16) CASE_T(32)
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
// Multi-rank: pack+exchange on the comm stream, interior rows meanwhile,
// boundary rows after the ghosts landed
// (hypre_ParCSRMatrixMatvecOutOfPlaceHost, src/parcsr_mv/par_csr_matvec.c:163-218).
// red_out: device scalars receiving the Epi::NRED fused reductions (local sums).
template <class Epi>
static int run_rows(b200_parcsr *A, const DMat &D, const DMat &O, bool use_halo,
                    const double *x, const Epi &epi, double *red_out)
{
   b200_ctx *ctx = A->ctx;
   if (use_halo && (A->num_sends + A->num_recvs) > 0)
   {
      B200_TRY(halo_start(A, x));
      B200_TRY((launch_rows_layout<true, Epi>(ctx, D, O.rp, x, epi, make_redctx(ctx, red_out, 0))));
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
   return launch_rows_layout<false, Epi>(ctx, D, nullptr, x, epi, make_redctx(ctx, red_out, 0));
}
