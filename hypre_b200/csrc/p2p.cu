// p2p.cu -- arena bootstrap (CUDA IPC), exchange channels, put kernel, small all-reduce.
#include "p2p.cuh"
#include <algorithm>

// ---------------------------------------------------------------- arena
static size_t arena_default_bytes()
{
   const char *e = getenv("B200_ARENA_MB");
   size_t mb = e ? (size_t)atoll(e) : 256;
   if (mb < 8) { mb = 8; }
   return mb << 20;
}

// bump allocator inside this rank's arena; 256-byte aligned, zero initialised at bootstrap
static int arena_alloc(b200_ctx *c, size_t bytes, size_t *off)
{
   size_t o = (c->arena_used + 255) & ~(size_t)255;
   if (o + bytes > c->arena_bytes)
   {
      b200_set_error("peer arena exhausted (%zu + %zu > %zu bytes): raise B200_ARENA_MB", o, bytes, c->arena_bytes);
      return B200_ERR_CUDA;
   }
   *off = o;
   c->arena_used = o + bytes;
   return B200_OK;
}

constexpr int AR_MAX = 256;     // doubles per small all-reduce (dots, coarse rhs)
constexpr int AR_BIG_MAX = 1 << 18;   // doubles per vector all-reduce (restriction onto a replicated level)
constexpr int AR_BIG_GRID = 2 * B200_NUM_SMS;   // all blocks must be co-resident (they wait on one another)

int p2p_bootstrap(b200_ctx *c)
{
   c->use_p2p = 0;
   if (c->nranks <= 1) { return B200_OK; }
   const char *mode = getenv("B200_HALO");
   if (mode && strcmp(mode, "nccl") == 0) { return B200_OK; }      // A/B switch: NCCL baseline
   c->arena_bytes = arena_default_bytes();
   CUDA_TRY(cudaMalloc((void **)&c->arena, c->arena_bytes));
   CUDA_TRY(cudaMemset(c->arena, 0, c->arena_bytes));
   c->arena_used = 0;
   CUDA_TRY(cudaMalloc((void **)&c->d_err, sizeof(int)));
   CUDA_TRY(cudaMemset(c->d_err, 0, sizeof(int)));

   // exchange the IPC handles through the NCCL communicator (the only host-level channel we have)
   cudaIpcMemHandle_t mine;
   CUDA_TRY(cudaIpcGetMemHandle(&mine, c->arena));
   char *d_h = nullptr;
   const size_t hs = sizeof(cudaIpcMemHandle_t);
   CUDA_TRY(cudaMalloc((void **)&d_h, hs * (size_t)(c->nranks + 1)));
   CUDA_TRY(cudaMemcpy(d_h + hs * c->nranks, &mine, hs, cudaMemcpyHostToDevice));
   NCCL_TRY(ncclAllGather(d_h + hs * c->nranks, d_h, hs, ncclChar, c->comm, c->stream));
   CUDA_TRY(cudaStreamSynchronize(c->stream));
   std::vector<cudaIpcMemHandle_t> all((size_t)c->nranks);
   CUDA_TRY(cudaMemcpy(all.data(), d_h, hs * c->nranks, cudaMemcpyDeviceToHost));
   CUDA_TRY(cudaFree(d_h));

   c->peer_arena.assign((size_t)c->nranks, nullptr);
   int ok = 1;
   for (int r = 0; r < c->nranks; r++)
   {
      if (r == c->rank) { c->peer_arena[r] = c->arena; continue; }
      void *p = nullptr;
      cudaError_t e = cudaIpcOpenMemHandle(&p, all[r], cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) { ok = 0; cudaGetLastError(); fprintf(stderr, "b200: cudaIpcOpenMemHandle(rank %d) failed: %s\n", r, cudaGetErrorString(e)); break; }
      c->peer_arena[r] = (char *)p;
   }
   // all ranks must agree (min over ranks) or nobody uses the peer path
   int *d_ok = nullptr;
   CUDA_TRY(cudaMalloc((void **)&d_ok, sizeof(int)));
   CUDA_TRY(cudaMemcpy(d_ok, &ok, sizeof(int), cudaMemcpyHostToDevice));
   NCCL_TRY(ncclAllReduce(d_ok, d_ok, 1, ncclInt, ncclMin, c->comm, c->stream));
   CUDA_TRY(cudaStreamSynchronize(c->stream));
   CUDA_TRY(cudaMemcpy(&ok, d_ok, sizeof(int), cudaMemcpyDeviceToHost));
   CUDA_TRY(cudaFree(d_ok));
   if (!ok)
   {
      fprintf(stderr, "b200: peer-memory halo path unavailable, using NCCL send/recv\n");
      return B200_OK;
   }
   // all-reduce slots: vals[2][nranks][AR_MAX] values with their sequence numbers (LLVal)
   B200_TRY(arena_alloc(c, sizeof(LLVal) * 2 * (size_t)c->nranks * AR_MAX, &c->ar_vals_off));
   B200_TRY(arena_alloc(c, sizeof(LLVal) * 2 * (size_t)c->nranks * AR_BIG_MAX, &c->arb_vals_off));
   CUDA_TRY(cudaMalloc((void **)&c->d_ar_seq, 4 * sizeof(u64)));     // [0] small seq, [1] big seq, [2..3] big tickets
   CUDA_TRY(cudaMemset(c->d_ar_seq, 0, 4 * sizeof(u64)));
   B200_TRY(dev_upload(&c->d_peer_arena, c->peer_arena.data(), c->peer_arena.size()));
   { const char *e = getenv("B200_P2P_DEBUG"); c->p2p_debug = e ? atoi(e) : 0; }
   c->use_p2p = 1;
   return B200_OK;
}

void p2p_shutdown(b200_ctx *c)
{
   if (!c->arena) { return; }
   for (int r = 0; r < (int)c->peer_arena.size(); r++)
   {
      if (r != c->rank && c->peer_arena[r]) { cudaIpcCloseMemHandle(c->peer_arena[r]); }
   }
   cudaFree(c->d_peer_arena); cudaFree(c->d_ar_seq); cudaFree(c->d_err);
   cudaFree(c->arena);
   c->arena = nullptr; c->use_p2p = 0;
}

// ---------------------------------------------------------------- channels
struct b200_p2p_chan {
   b200_ctx *ctx = nullptr;
   P2PChan   v;
   int      *d_out_starts = nullptr;
   LLVal   **d_out_dst[2] = {nullptr, nullptr};
   u64     **d_in_ack = nullptr;
   int      *d_out_wait_ack = nullptr, *d_in_send_ack = nullptr;
   u64      *d_seq = nullptr;       // [2] seq_put, seq_wait
   unsigned *d_ticket = nullptr;    // [2]
};

const P2PChan &p2p_chan_view(const b200_p2p_chan *ch) { return ch->v; }

void p2p_chan_destroy(b200_p2p_chan *ch)
{
   if (!ch) { return; }
   cudaFree(ch->d_out_starts); cudaFree(ch->d_out_dst[0]); cudaFree(ch->d_out_dst[1]);
   cudaFree(ch->d_out_wait_ack); cudaFree(ch->d_in_send_ack); cudaFree(ch->d_in_ack); cudaFree(ch->d_seq); cudaFree(ch->d_ticket);
   delete ch;
}

// Small host-level neighbour exchange of `w` 64-bit words per message over the halo communicator:
// msg_out[i*w..] goes to out_procs[i], msg_in[j*w..] comes from in_procs[j].
static int exchange_words(b200_ctx *c, int w, int n_out, const int *out_procs, const long long *msg_out,
                          int n_in, const int *in_procs, long long *msg_in)
{
   long long *d_o = nullptr, *d_i = nullptr;
   CUDA_TRY(cudaMalloc((void **)&d_o, sizeof(long long) * (size_t)(n_out * w + 1)));
   CUDA_TRY(cudaMalloc((void **)&d_i, sizeof(long long) * (size_t)(n_in * w + 1)));
   if (n_out > 0) { CUDA_TRY(cudaMemcpy(d_o, msg_out, sizeof(long long) * (size_t)(n_out * w), cudaMemcpyHostToDevice)); }
   NCCL_TRY(ncclGroupStart());
   for (int j = 0; j < n_in; j++)  { NCCL_TRY(ncclRecv(d_i + (size_t)j * w, (size_t)w, ncclInt64, in_procs[j], c->hcomm, c->cstream)); }
   for (int i = 0; i < n_out; i++) { NCCL_TRY(ncclSend(d_o + (size_t)i * w, (size_t)w, ncclInt64, out_procs[i], c->hcomm, c->cstream)); }
   NCCL_TRY(ncclGroupEnd());
   CUDA_TRY(cudaStreamSynchronize(c->cstream));
   if (n_in > 0) { CUDA_TRY(cudaMemcpy(msg_in, d_i, sizeof(long long) * (size_t)(n_in * w), cudaMemcpyDeviceToHost)); }
   CUDA_TRY(cudaFree(d_o)); CUDA_TRY(cudaFree(d_i));
   return B200_OK;
}

int p2p_chan_create(b200_ctx *c, int n_out, const int *out_procs, const int *out_starts,
                    int n_in, const int *in_procs, const int *in_starts, b200_p2p_chan **out)
{
   *out = nullptr;
   if (!c->use_p2p) { return B200_OK; }
   b200_p2p_chan *ch = new b200_p2p_chan();
   ch->ctx = c;
   const int in_total = n_in > 0 ? in_starts[n_in] : 0;
   size_t off_buf[2] = {0, 0}, off_ack = 0;
   int s = B200_OK;
   for (int p = 0; p < 2 && s == B200_OK; p++) { s = arena_alloc(c, sizeof(LLVal) * (size_t)(in_total + 1), &off_buf[p]); }
   if (s == B200_OK) { s = arena_alloc(c, sizeof(u64) * (size_t)(n_out + 1), &off_ack); }
   // A rank whose arena is exhausted still takes part in the two exchanges below -- its neighbours are already
   // inside them -- and sends offsets of -1, so that every neighbour reports the failure instead of waiting for it.
   const int  arena_status = s;
   const bool arena_ok = (s == B200_OK);

   // tell every sender where its segment lands in my arena (2 words + 1 spare), and every receiver
   // where its acknowledgement lands (1 word)
   std::vector<long long> to_in((size_t)n_in * 3 + 1), from_out((size_t)n_out * 3 + 1);
   for (int j = 0; j < n_in; j++)
   {
      to_in[(size_t)j * 3 + 0] = arena_ok ? (long long)(off_buf[0] + sizeof(LLVal) * (size_t)in_starts[j]) : -1LL;
      to_in[(size_t)j * 3 + 1] = arena_ok ? (long long)(off_buf[1] + sizeof(LLVal) * (size_t)in_starts[j]) : -1LL;
      to_in[(size_t)j * 3 + 2] = 0;
   }
   s = exchange_words(c, 3, n_in, in_procs, to_in.data(), n_out, out_procs, from_out.data());
   std::vector<long long> to_out((size_t)n_out + 1), from_in((size_t)n_in + 1);
   for (int i = 0; i < n_out; i++) { to_out[(size_t)i] = arena_ok ? (long long)(off_ack + sizeof(u64) * (size_t)i) : -1LL; }
   if (s == B200_OK) { s = exchange_words(c, 1, n_out, out_procs, to_out.data(), n_in, in_procs, from_in.data()); }
   if (!arena_ok) { delete ch; return arena_status; }
   if (s != B200_OK) { delete ch; return s; }
   for (int i = 0; i < n_out; i++)
   {
      if (from_out[(size_t)i * 3 + 0] < 0 || from_out[(size_t)i * 3 + 1] < 0)
      {
         b200_set_error("peer arena exhausted on rank %d (B200_ARENA_MB sizes it)", out_procs[i]);
         delete ch;
         return B200_ERR_CUDA;
      }
   }
   for (int j = 0; j < n_in; j++)
   {
      if (from_in[(size_t)j] < 0)
      {
         b200_set_error("peer arena exhausted on rank %d (B200_ARENA_MB sizes it)", in_procs[j]);
         delete ch;
         return B200_ERR_CUDA;
      }
   }

   std::vector<LLVal *> dst0((size_t)n_out + 1), dst1((size_t)n_out + 1);
   std::vector<u64 *> iack((size_t)n_in + 1);
   // per pair: a receiver I do not receive from must acknowledge (see p2p.cuh, flow control)
   std::vector<int> wait_ack((size_t)n_out + 1, 0), send_ack((size_t)n_in + 1, 0);
   int n_ack_in = 0;
   for (int i = 0; i < n_out; i++)
   {
      char *base = c->peer_arena[(size_t)out_procs[i]];
      dst0[(size_t)i]  = (LLVal *)(base + from_out[(size_t)i * 3 + 0]);
      dst1[(size_t)i]  = (LLVal *)(base + from_out[(size_t)i * 3 + 1]);
      wait_ack[(size_t)i] = std::find(in_procs, in_procs + n_in, out_procs[i]) == in_procs + n_in ? 1 : 0;
   }
   for (int j = 0; j < n_in; j++)
   {
      send_ack[(size_t)j] = std::find(out_procs, out_procs + n_out, in_procs[j]) == out_procs + n_out ? 1 : 0;
      n_ack_in += send_ack[(size_t)j];
   }
   for (int j = 0; j < n_in; j++) { iack[(size_t)j] = (u64 *)(c->peer_arena[(size_t)in_procs[j]] + from_in[(size_t)j]); }
   std::vector<int> zero_starts(1, 0);
   s = dev_upload(&ch->d_out_starts, n_out > 0 ? out_starts : zero_starts.data(), (size_t)n_out + 1);
   if (s == B200_OK) { s = dev_upload(&ch->d_out_dst[0], dst0.data(), dst0.size()); }
   if (s == B200_OK) { s = dev_upload(&ch->d_out_dst[1], dst1.data(), dst1.size()); }
   if (s == B200_OK) { s = dev_upload(&ch->d_out_wait_ack, wait_ack.data(), wait_ack.size()); }
   if (s == B200_OK) { s = dev_upload(&ch->d_in_send_ack, send_ack.data(), send_ack.size()); }
   if (s == B200_OK) { s = dev_upload(&ch->d_in_ack, iack.data(), iack.size()); }
   if (s == B200_OK && cudaMalloc((void **)&ch->d_seq, 2 * sizeof(u64)) != cudaSuccess) { s = B200_ERR_CUDA; }
   if (s == B200_OK && cudaMalloc((void **)&ch->d_ticket, 2 * sizeof(unsigned)) != cudaSuccess) { s = B200_ERR_CUDA; }
   if (s != B200_OK) { p2p_chan_destroy(ch); return s; }
   cudaMemset(ch->d_seq, 0, 2 * sizeof(u64));
   cudaMemset(ch->d_ticket, 0, 2 * sizeof(unsigned));

   P2PChan &v = ch->v;
   v.n_out = n_out; v.n_in = n_in;
   v.n_out_total = n_out > 0 ? out_starts[n_out] : 0;
   v.out_starts = ch->d_out_starts;
   v.out_dst[0] = ch->d_out_dst[0]; v.out_dst[1] = ch->d_out_dst[1];
   v.out_ack = (u64 *)(c->arena + off_ack);
   v.out_wait_ack = ch->d_out_wait_ack;
   v.in_buf[0] = (LLVal *)(c->arena + off_buf[0]); v.in_buf[1] = (LLVal *)(c->arena + off_buf[1]);
   v.in_ack = ch->d_in_ack;
   v.in_send_ack = ch->d_in_send_ack;
   v.n_ack_in = n_ack_in;
   v.seq_put = ch->d_seq; v.seq_wait = ch->d_seq + 1;
   v.ticket = ch->d_ticket;
   v.err = c->d_err;
   *out = ch;
   return B200_OK;
}

// ---------------------------------------------------------------- all-reduce
// Every rank stores its contribution into slot [parity][rank] of every peer's arena -- each value with the sequence
// number of this reduction (flag-in-data, p2p.cuh) -- and sums the slots of its own arena in rank order as they
// arrive: deterministic and bit-identical on all ranks, one launch, no fence, no library call.  Double buffered by
// the parity of the sequence number; a rank can start reduction s + 2 only after it finished s + 1, for which every
// peer had to have finished reading s.
__global__ void __launch_bounds__(256) p2p_allreduce_kernel(char *const *__restrict__ peer, size_t vals_off,
                                                            int rank, int nranks, u64 *seq_p, double *d, int cnt, int *err,
                                                            const int *stop)
{
   if (*stop != 0) { return; }      // identical on all ranks (derived from all-reduced scalars)
   const u64 seq = ld_relaxed_gpu(seq_p) + 1;
   const unsigned flag = (unsigned)seq;
   const int par = (int)(seq & 1);
   for (int k = threadIdx.x; k < cnt; k += blockDim.x)
   {
      const double v = d[k];
      for (int t = 0; t < nranks; t++)
      {
         ll_store((LLVal *)(peer[t] + vals_off) + ((size_t)par * nranks + rank) * AR_MAX + k, v, flag);
      }
   }
   const LLVal *mine = (const LLVal *)(peer[rank] + vals_off) + (size_t)par * nranks * AR_MAX;
   for (int k = threadIdx.x; k < cnt; k += blockDim.x)
   {
      double s = 0.0;
      for (int t = 0; t < nranks; t++) { s += ll_read(mine + (size_t)t * AR_MAX + k, flag, err); }
      d[k] = s;
   }
   __syncthreads();
   if (threadIdx.x == 0) { *seq_p = seq; }
}

int p2p_allreduce(b200_ctx *c, double *d, int cnt)
{
   if (cnt > AR_MAX) { b200_set_error("p2p_allreduce: %d > %d values", cnt, AR_MAX); return B200_ERR_ARG; }
   LAUNCH(c, p2p_allreduce_kernel, 1, 256, 0, c->d_peer_arena, c->ar_vals_off, c->rank, c->nranks,
          c->d_ar_seq, d, cnt, c->d_err, c->d_stop);
   CHECK_LAUNCH();
   return B200_OK;
}

// Vector all-reduce (n <= AR_BIG_MAX): same protocol with several blocks.  Block b owns the elements
// k = b*256 + tid (grid stride) in both phases; the last block to finish advances the sequence counter.
__global__ void __launch_bounds__(256) p2p_allreduce_big_kernel(char *const *__restrict__ peer, size_t vals_off,
                                                                int rank, int nranks, u64 *seq_p, unsigned *tickets,
                                                                double *d, int n, int *err, const int *stop)
{
   if (*stop != 0) { return; }
   const u64 seq = ld_relaxed_gpu(seq_p) + 1;
   const unsigned flag = (unsigned)seq;
   const int par = (int)(seq & 1);
   for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x)
   {
      const double v = d[k];
      for (int t = 0; t < nranks; t++)
      {
         ll_store((LLVal *)(peer[t] + vals_off) + ((size_t)par * nranks + rank) * AR_BIG_MAX + k, v, flag);
      }
   }
   const LLVal *mine = (const LLVal *)(peer[rank] + vals_off) + (size_t)par * nranks * AR_BIG_MAX;
   for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x)
   {
      double s = 0.0;
      for (int t = 0; t < nranks; t++) { s += ll_read(mine + (size_t)t * AR_BIG_MAX + k, flag, err); }
      d[k] = s;
   }
   __syncthreads();
   if (threadIdx.x == 0)
   {
      // every block read the counter before it took its ticket: the last one advances it
      __threadfence();
      if (atomicAdd(tickets + 1, 1u) == gridDim.x - 1) { *seq_p = seq; tickets[1] = 0u; __threadfence(); }
   }
}

int p2p_allreduce_vec_max(void) { return AR_BIG_MAX; }

int p2p_allreduce_vec(b200_ctx *c, double *d, int n)
{
   if (n > AR_BIG_MAX) { b200_set_error("p2p_allreduce_vec: %d > %d values", n, AR_BIG_MAX); return B200_ERR_ARG; }
   if (n <= 0) { return B200_OK; }
   int grid = (n + 255) / 256;
   if (grid > AR_BIG_GRID) { grid = AR_BIG_GRID; }
   LAUNCH(c, p2p_allreduce_big_kernel, grid, 256, 0, c->d_peer_arena, c->arb_vals_off, c->rank, c->nranks,
          c->d_ar_seq + 1, (unsigned *)(c->d_ar_seq + 2), d, n, c->d_err, c->d_stop);
   CHECK_LAUNCH();
   return B200_OK;
}

int ctx_rendezvous(b200_ctx *c)
{
   if (c->nranks <= 1) { return B200_OK; }
   NCCL_TRY(ncclAllReduce(c->d_scalars + (B200_NUM_SCALARS - 1), c->d_scalars + (B200_NUM_SCALARS - 1), 1, ncclDouble,
                          ncclSum, c->comm, c->stream));
   return B200_OK;
}
