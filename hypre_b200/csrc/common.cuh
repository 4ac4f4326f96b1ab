// common.cuh -- internal types of libb200amg (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <nccl.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <string>
#include <type_traits>
#include "../../include/b200amg.h"

// ---------------------------------------------------------------- errors
void b200_set_error(const char *fmt, ...);

#define CUDA_TRY(expr)                                                              \
   do {                                                                             \
      cudaError_t _e = (expr);                                                      \
      if (_e != cudaSuccess) {                                                      \
         b200_set_error("CUDA error %s at %s:%d: %s", cudaGetErrorName(_e),         \
                        __FILE__, __LINE__, cudaGetErrorString(_e));                \
         return B200_ERR_CUDA;                                                      \
      }                                                                             \
   } while (0)

#define NCCL_TRY(expr)                                                              \
   do {                                                                             \
      ncclResult_t _r = (expr);                                                     \
      if (_r != ncclSuccess) {                                                      \
         b200_set_error("NCCL error at %s:%d: %s", __FILE__, __LINE__,              \
                        ncclGetErrorString(_r));                                    \
         return B200_ERR_NCCL;                                                      \
      }                                                                             \
   } while (0)

#define B200_TRY(expr)                                                              \
   do { int _s = (expr); if (_s != B200_OK) return _s; } while (0)

#define ARG_CHECK(cond)                                                             \
   do {                                                                             \
      if (!(cond)) {                                                                \
         b200_set_error("bad argument: %s (%s:%d)", #cond, __FILE__, __LINE__);     \
         return B200_ERR_ARG;                                                       \
      }                                                                             \
   } while (0)

// ---------------------------------------------------------------- context
struct ProfSample { const char *name; cudaEvent_t e0, e1; };
constexpr int B200_NUM_SMS      = 148;   // B200: 2 dies x 74 SMs
constexpr int B200_MAX_RED      = 4;     // fused reductions per kernel
constexpr int B200_MAX_PARTIALS = 8192;  // per-block partial sums kept per reduction slot

struct b200_ctx {
   int          device   = 0;
   int          rank     = 0;
   int          nranks   = 1;
   cudaStream_t stream   = nullptr;   // compute stream (all kernels)
   cudaStream_t cstream  = nullptr;   // halo / collective stream
   cudaEvent_t  ev_pack  = nullptr;   // pack done   (compute -> comm)
   cudaEvent_t  ev_halo  = nullptr;   // halo landed (comm -> compute)
   cudaEvent_t  ev_t0 = nullptr, ev_t1 = nullptr;
   ncclComm_t   comm     = nullptr;   // collectives (dots, coarse gather) on `stream`
   ncclComm_t   hcomm    = nullptr;   // halo point-to-point on `cstream` (own communicator: the two
                                      // streams run concurrently and must not share one)
   long long    launches = 0;
   void        *tail = nullptr;      // recorder / program cache of the persistent cycle kernel (tail.cuh)
   long long    generation = 0;      // bumped whenever device objects baked into a captured graph go away or change
   // peer-memory exchange (p2p.cuh): this rank's arena, the IPC-mapped arenas of all ranks
   int          use_p2p = 0;
   int          p2p_debug = 0;       // B200_P2P_DEBUG (timing experiments)
   char        *arena = nullptr;
   size_t       arena_bytes = 0, arena_used = 0;
   std::vector<char *> peer_arena;
   char       **d_peer_arena = nullptr;
   size_t       ar_vals_off = 0, ar_flag_off = 0, arb_vals_off = 0, arb_flag_off = 0;
   unsigned long long *d_ar_seq = nullptr;
   int         *d_err = nullptr;      // set by a kernel whose bounded spin expired
   int         *h_err = nullptr;      // pinned mirror
   // reduction scratch: partials[slot][block], finished-block counter, results
   double      *d_partials = nullptr;
   unsigned    *d_counter  = nullptr;
   double      *d_scalars  = nullptr;  // device-resident scalars (dots, alpha, beta, ...)
   double      *h_scalars  = nullptr;  // pinned mirror
   // L2 flush scratch
   char        *d_flush = nullptr;
   size_t       flush_bytes = 0;
   // pinned staging for host-vector entry points
   double      *h_stage = nullptr;
   size_t       h_stage_n = 0;
   bool         capturing = false;
   int          prof = 0;            // B200_PROFILE
   std::vector<ProfSample> prof_samples;
   int         *d_stop = nullptr;     // device stop word checked by every kernel (see RedCtx::stop)
   int         *d_iter = nullptr;     // device iteration counter of the running PCG solve
   void        *h_ring = nullptr;     // mapped pinned ring of per-iteration records (pcg.cu)
   void        *d_ring = nullptr;     // device alias of h_ring
   long long    pcg_epoch = 0;
   cudaGraphExec_t pcg_graph = nullptr;
   const void  *pcg_graph_key[8] = {nullptr};
   long long    pcg_graph_launches = 0;
   // cached device work vectors (PCG p/s/r, host-entry b/x, ...): grown on demand, never
   // freed before the context dies, so repeated solves do not pay cudaMalloc/cudaFree
   static constexpr int NUM_WS = 12;
   double      *ws[NUM_WS] = {nullptr};
   size_t       ws_n[NUM_WS] = {0};
   // pinned mirror of the device scalars of a GMRES / BiCGSTAB solve (krylov.cu)
   double      *h_kry = nullptr;
   size_t       h_kry_n = 0;
};

enum { WS_PCG_P = 0, WS_PCG_S, WS_PCG_R, WS_HOST_B, WS_HOST_X, WS_HOST_Y, WS_AMG_R, WS_KRYLOV, WS_PCG_P2, WS_PCG_Z };
// device work vector `slot` with at least n doubles
int ctx_workspace(b200_ctx *c, int slot, size_t n, double **out);

constexpr int B200_NUM_SCALARS = 64;

// B200_PROFILE=1: every launch is bracketed by CUDA events and accounted per kernel name
// (b200_prof_report); used to see where a multi-rank step spends its time, where ncu cannot go
void prof_begin(b200_ctx *c, const char *name);
void prof_end(b200_ctx *c);

#define LAUNCH(ctx, kern, grid, block, smem, ...)                                   \
   do {                                                                             \
      if ((ctx)->prof) { prof_begin((ctx), #kern); }                                \
      kern<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);                \
      if ((ctx)->prof) { prof_end((ctx)); }                                         \
      (ctx)->launches++;                                                            \
   } while (0)

#define CHECK_LAUNCH()  CUDA_TRY(cudaGetLastError())

// ---------------------------------------------------------------- matrices
// One CSR block in HBM.  Three layouts:
//   layout 0 (CSR-vector): rp/ci/va as given; `lanes` lanes of a warp cooperate on a row
//   layout 2 (sliced ELL with row sorting, "SELL-32-256"): as layout 1, but inside every window of
//            256 rows the rows are stably sorted by length before they are cut into slices, so
//            ragged rows (interpolation: 1-4 entries, coarse operators) pad almost nothing;
//            perm[p] = offset in its window of the row stored at position p
//   layout 1 (sliced ELL, slice height 32): slice s holds rows 32s..32s+31, entry k of
//            lane l at sl_off[s] + k*32 + l  -> every warp load is one contiguous
//            128 B (ci) / 256 B (va) segment; padded entries have va = 0.
struct DMat {
   int        n_rows = 0, n_cols = 0;
   long long  nnz = 0;
   int       *rp = nullptr;    // [n_rows+1]
   int       *ci = nullptr;    // [nnz]
   double    *va = nullptr;    // [nnz]
   int        layout = 0;
   int        lanes  = 4;
   int        n_slices = 0;
   long long  sell_entries = 0;
   long long *sl_off = nullptr;  // [n_slices+1]
   unsigned short *sl_perm = nullptr;   // layout 2: [n_slices*32]
   std::vector<int> h_pos;              // layout 2: storage position of every row (host)
   int       *sl_ci  = nullptr;
   double    *sl_va  = nullptr;
   int        max_row = 0;
   double     avg_row = 0.0;
   // compressed thread-per-row storage (cfmt != 0, layouts 1/2 only): 16-bit column offsets from a
   // base per (slice, entry), values through a dictionary when the block holds few distinct ones
   int        cfmt = 0;                  // CFMT_*
   unsigned short *c_idx  = nullptr;     // [sell_entries] chunk-transposed: slice, chunk of 4 entries, lane, entry
   int            *c_base = nullptr;     // [sell_entries/32] base column of (slice, entry) or CFMT_ESC
   void           *c_code = nullptr;     // [sell_entries] value codes (u8 / u16), chunk-transposed like c_idx
   double         *c_dict = nullptr;     // distinct values
   int             n_dict = 0;
   long long       stored_bytes = 0;     // bytes one pass over the block reads from the matrix arrays
   int             zpad = 0;             // compressed storage without CFMT_PAD lanes (padding = valid column x value 0)
   int             pat_stride = 0;       // CFMT_PAT8: 16-byte units per slice header (headers in c_code, lane words in c_idx)
};

// cfmt: value encoding of the compressed storage (column indices are always base + u16)
// CFMT_PAT8: one byte per entry = index of its (column - row, value) pattern, lane-uniform entries once per slice
enum { CFMT_NONE = 0, CFMT_VAL8 = 1, CFMT_VAL16 = 2, CFMT_VAL64 = 3, CFMT_PAT8 = 4 };
constexpr int            CFMT_ESC = (int)0x80000000;   // c_base sentinel: this (slice, entry) keeps 32-bit columns in sl_ci
constexpr unsigned short CFMT_PAD = 0xFFFFu;           // c_idx value of a padding lane

// view passed by value to kernels
struct MatView {
   int              n_rows;
   const int       *rp;
   const int       *ci;
   const double    *va;
   const long long *sl_off;
   const unsigned short *perm;   // layout 2 only, else nullptr
   const int       *sl_ci;
   const double    *sl_va;
   const unsigned short *c_idx;  // compressed storage (DMat::cfmt != 0), else nullptr
   const int       *c_base;
   const void      *c_code;
   const double    *c_dict;
   int              n_dict;      // entries of c_dict (<= 256 for CFMT_VAL8)
   int              aux;         // CFMT_PAT8: 16-byte units per slice header
};

static inline MatView view_of(const DMat &m)
{
   MatView v;
   v.n_rows = m.n_rows; v.rp = m.rp; v.ci = m.ci; v.va = m.va;
   v.sl_off = m.sl_off; v.sl_ci = m.sl_ci; v.sl_va = m.sl_va; v.perm = m.sl_perm;
   v.c_idx = m.c_idx; v.c_base = m.c_base; v.c_code = m.c_code; v.c_dict = m.c_dict; v.n_dict = m.n_dict; v.aux = m.pat_stride;
   return v;
}

struct HostCsr {
   int n_rows = 0, n_cols = 0;
   std::vector<int>    i, j;
   std::vector<double> a;
};

struct b200_p2p_chan;

struct b200_parcsr {
   b200_ctx *ctx = nullptr;
   // peer-memory channels (N > 1): forward = ghost values of x (job 1), reverse = contributions
   // to the owners for A^T x (job 2); nullptr -> NCCL send/recv path
   b200_p2p_chan *ch_fwd = nullptr, *ch_rev = nullptr;
   int n_rows = 0, n_cols_diag = 0, n_cols_offd = 0;
   DMat diag, offd;
   // explicit transposes (restriction) -- built on first MatvecT
   bool hasT = false;
   DMat diagT, offdT;
   // strict lower triangle of diag + the diagonal (two-stage GS) -- built on demand
   bool hasL = false;
   DMat lower;
   double *d_diagonal = nullptr;
   bool zero_diagonal = false;
   std::vector<int> gs_block_of;     // gathered level: owner block of every row (two-stage GS stays block-local)
   // host copies kept for derived operators
   HostCsr h_diag, h_offd;
   // halo plan (hypre_ParCSRCommPkg)
   int num_sends = 0, num_recvs = 0;
   std::vector<int> send_procs, send_map_starts, recv_procs, recv_vec_starts;
   int     n_send_elmts = 0;
   int    *d_send_map  = nullptr;  // [n_send_elmts] local rows to pack
   double *d_send_buf  = nullptr;  // [n_send_elmts]
   double *d_ghost     = nullptr;  // [n_cols_offd]   x_tmp of the reference
   // rows with at least one offd entry (processed after the halo landed)
   int     n_bnd = 0;
   int    *d_bnd_rows = nullptr;
   // for MatvecT (job 2): contributions for ghost columns / received contributions
   double *d_ghost_out = nullptr;  // [n_cols_offd]
   double *d_recv_buf  = nullptr;  // [n_send_elmts]
   // deterministic unpack plan for job 2: CSR of (local row -> positions in recv buf)
   int    *d_unpack_rp = nullptr, *d_unpack_pos = nullptr;
   unsigned char *d_grp_fwd = nullptr;    // [ceil(n_rows/32)] group has a row with offd entries
   unsigned char *d_grp_rev = nullptr;
   int     int_fwd[2] = {0, 0}, int_rev[2] = {0, 0};   // longest runs of groups without remote data    // [ceil(n_cols_diag/32)] group has a row receiving A^T contributions
   int    *d_unpack_rp_dense = nullptr;   // [n_cols_diag+1]: the same lists indexed by local row (fused kernel)
   int     n_unpack_rows = 0;
   int    *d_unpack_rows = nullptr;
};

// ---------------------------------------------------------------- helpers
template <class T>
static inline int dev_upload(T **dst, const T *src, size_t n)
{
   *dst = nullptr;
   if (n == 0) { return B200_OK; }
   CUDA_TRY(cudaMalloc((void **)dst, n * sizeof(T)));
   CUDA_TRY(cudaMemcpy(*dst, src, n * sizeof(T), cudaMemcpyHostToDevice));
   return B200_OK;
}

// host image of the layout arrays of a DMat (pack.cu: pure host code, CPU-tested)
struct PackedDMat {
   int layout = 0, lanes = 4, n_slices = 0, max_row = 0, cfmt = CFMT_NONE;
   double avg_row = 0.0;
   long long padded = 0, stored_bytes = 0, n_esc = 0;
   int zpad = 0;
   std::vector<long long>      sl_off;      // [n_slices+1]
   std::vector<int>            sci;         // entry-major 32-bit columns (plain layout; escape groups)
   std::vector<double>         sva;         // entry-major fp64 values (plain layout; CFMT_VAL64)
   std::vector<unsigned short> perm;        // layout 2
   std::vector<int>            h_pos;       // layout 2
   std::vector<unsigned short> idx;         // chunk-transposed 16-bit column offsets
   std::vector<int>            base;        // [padded/32]
   std::vector<unsigned char>  code;        // chunk-transposed value codes (u8, or u16 as byte pairs)
   std::vector<double>         dict;        // value dictionary; CFMT_PAT8: (value, delta bits) pairs
   std::vector<unsigned char>  pat;         // CFMT_PAT8: slice headers, pat_stride x 16 bytes each
   std::vector<unsigned>       pat_lanes;   // CFMT_PAT8: lane words (32 per header word that differs between lanes)
   int                         pat_stride = 0;
};
void dmat_pack(const HostCsr &h, int allow_sell, PackedDMat *P);
int  packed_matvec_host(const PackedDMat &P, const HostCsr &h, const double *x, double *y);
int  dmat_upload(b200_ctx *ctx, const HostCsr &h, DMat *m, int allow_sell);
void dmat_free(DMat *m);
void host_transpose(const HostCsr &a, HostCsr *t);

// internal launchers shared between translation units (all on ctx->stream)
int halo_start(b200_parcsr *A, const double *x);   // pack + post exchange on cstream
int halo_finish(b200_parcsr *A);                   // make compute stream wait for ghosts
