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
   // all-reduce slots: vals[2][nranks][AR_MAX] doubles, flag[nranks] u64, then a local seq word
   B200_TRY(arena_alloc(c, sizeof(double) * 2 * (size_t)c->nranks * AR_MAX, &c->ar_vals_off));
   B200_TRY(arena_alloc(c, sizeof(u64) * (size_t)c->nranks, &c->ar_flag_off));
   B200_TRY(arena_alloc(c, sizeof(double) * 2 * (size_t)c->nranks * AR_BIG_MAX, &c->arb_vals_off));
   B200_TRY(arena_alloc(c, sizeof(u64) * (size_t)c->nranks, &c->arb_flag_off));
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
   double  **d_out_dst[2] = {nullptr, nullptr};
   u64     **d_out_flag = nullptr;
   u64     **d_in_ack = nullptr;
   u64      *d_seq = nullptr;       // [2] seq_put, seq_wait
   unsigned *d_ticket = nullptr;    // [2]
};

const P2PChan &p2p_chan_view(const b200_p2p_chan *ch) { return ch->v; }

void p2p_chan_destroy(b200_p2p_chan *ch)
{
   if (!ch) { return; }
   cudaFree(ch->d_out_starts); cudaFree(ch->d_out_dst[0]); cudaFree(ch->d_out_dst[1]);
   cudaFree(ch->d_out_flag); cudaFree(ch->d_in_ack); cudaFree(ch->d_seq); cudaFree(ch->d_ticket);
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
   size_t off_buf[2] = {0, 0}, off_flag = 0, off_ack = 0;
   int s = B200_OK;
   for (int p = 0; p < 2 && s == B200_OK; p++) { s = arena_alloc(c, sizeof(double) * (size_t)(in_total + 1), &off_buf[p]); }
   if (s == B200_OK) { s = arena_alloc(c, sizeof(u64) * (size_t)(n_in + 1), &off_flag); }
   if (s == B200_OK) { s = arena_alloc(c, sizeof(u64) * (size_t)(n_out + 1), &off_ack); }
   if (s != B200_OK) { delete ch; return s; }

   // tell every sender where its segment lands in my arena (3 words), and every receiver
   // where its acknowledgement lands (1 word)
   std::vector<long long> to_in((size_t)n_in * 3 + 1), from_out((size_t)n_out * 3 + 1);
   for (int j = 0; j < n_in; j++)
   {
      to_in[(size_t)j * 3 + 0] = (long long)(off_buf[0] + sizeof(double) * (size_t)in_starts[j]);
      to_in[(size_t)j * 3 + 1] = (long long)(off_buf[1] + sizeof(double) * (size_t)in_starts[j]);
      to_in[(size_t)j * 3 + 2] = (long long)(off_flag + sizeof(u64) * (size_t)j);
   }
   s = exchange_words(c, 3, n_in, in_procs, to_in.data(), n_out, out_procs, from_out.data());
   std::vector<long long> to_out((size_t)n_out + 1), from_in((size_t)n_in + 1);
   for (int i = 0; i < n_out; i++) { to_out[(size_t)i] = (long long)(off_ack + sizeof(u64) * (size_t)i); }
   if (s == B200_OK) { s = exchange_words(c, 1, n_out, out_procs, to_out.data(), n_in, in_procs, from_in.data()); }
   if (s != B200_OK) { delete ch; return s; }

   std::vector<double *> dst0((size_t)n_out + 1), dst1((size_t)n_out + 1);
   std::vector<u64 *> oflag((size_t)n_out + 1), iack((size_t)n_in + 1);
   for (int i = 0; i < n_out; i++)
   {
      char *base = c->peer_arena[(size_t)out_procs[i]];
      dst0[(size_t)i]  = (double *)(base + from_out[(size_t)i * 3 + 0]);
      dst1[(size_t)i]  = (double *)(base + from_out[(size_t)i * 3 + 1]);
      oflag[(size_t)i] = (u64 *)(base + from_out[(size_t)i * 3 + 2]);
   }
   for (int j = 0; j < n_in; j++) { iack[(size_t)j] = (u64 *)(c->peer_arena[(size_t)in_procs[j]] + from_in[(size_t)j]); }
   std::vector<int> zero_starts(1, 0);
   s = dev_upload(&ch->d_out_starts, n_out > 0 ? out_starts : zero_starts.data(), (size_t)n_out + 1);
   if (s == B200_OK) { s = dev_upload(&ch->d_out_dst[0], dst0.data(), dst0.size()); }
   if (s == B200_OK) { s = dev_upload(&ch->d_out_dst[1], dst1.data(), dst1.size()); }
   if (s == B200_OK) { s = dev_upload(&ch->d_out_flag, oflag.data(), oflag.size()); }
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
   v.out_flag = ch->d_out_flag;
   v.out_ack = (u64 *)(c->arena + off_ack);
   v.in_buf[0] = (double *)(c->arena + off_buf[0]); v.in_buf[1] = (double *)(c->arena + off_buf[1]);
   v.in_flag = (u64 *)(c->arena + off_flag);
   v.in_ack = ch->d_in_ack;
   v.seq_put = ch->d_seq; v.seq_wait = ch->d_seq + 1;
   v.ticket = ch->d_ticket;
   v.err = c->d_err;
   *out = ch;
   return B200_OK;
}

// ---------------------------------------------------------------- all-reduce
// Every rank stores its contribution into slot [parity][rank] of every peer's arena, publishes
// the sequence number, waits for all peers and sums the slots in rank order: deterministic and
// bit-identical on all ranks.  One launch, no library call.
__global__ void __launch_bounds__(256) p2p_allreduce_kernel(char *const *__restrict__ peer, size_t vals_off, size_t flag_off,
                                                            int rank, int nranks, u64 *seq_p, double *d, int cnt, int *err,
                                                            const int *stop)
{
   if (*stop != 0) { return; }      // identical on all ranks (derived from all-reduced scalars)
   const u64 seq = ld_relaxed_gpu(seq_p) + 1;
   const int par = (int)(seq & 1);
   for (int t = 0; t < nranks; t++)
   {
      double *slot = (double *)(peer[t] + vals_off) + ((size_t)par * nranks + rank) * AR_MAX;
      for (int k = threadIdx.x; k < cnt; k += blockDim.x) { slot[k] = d[k]; }
   }
   __threadfence_system();
   __syncthreads();
   if (threadIdx.x < nranks) { st_release_sys((u64 *)(peer[threadIdx.x] + flag_off) + rank, seq); }
   if (threadIdx.x < nranks) { p2p_spin_ge((const u64 *)(peer[rank] + flag_off) + threadIdx.x, seq, err); }
   __syncthreads();
   const double *mine = (const double *)(peer[rank] + vals_off) + (size_t)par * nranks * AR_MAX;
   for (int k = threadIdx.x; k < cnt; k += blockDim.x)
   {
      double s = 0.0;
      for (int t = 0; t < nranks; t++) { s += __ldcg(mine + (size_t)t * AR_MAX + k); }
      d[k] = s;
   }
   __syncthreads();
   if (threadIdx.x == 0) { *seq_p = seq; }
}

int p2p_allreduce(b200_ctx *c, double *d, int cnt)
{
   if (cnt > AR_MAX) { b200_set_error("p2p_allreduce: %d > %d values", cnt, AR_MAX); return B200_ERR_ARG; }
   LAUNCH(c, p2p_allreduce_kernel, 1, 256, 0, c->d_peer_arena, c->ar_vals_off, c->ar_flag_off, c->rank, c->nranks,
          c->d_ar_seq, d, cnt, c->d_err, c->d_stop);
   CHECK_LAUNCH();
   return B200_OK;
}

// Vector all-reduce (n <= AR_BIG_MAX): same protocol with several blocks.  Block b owns the elements
// k = b*256 + tid (grid stride) in both phases; the last block to finish its stores publishes the
// sequence number, every block then waits for all ranks and sums its elements in rank order.
__global__ void __launch_bounds__(256) p2p_allreduce_big_kernel(char *const *__restrict__ peer, size_t vals_off, size_t flag_off,
                                                                int rank, int nranks, u64 *seq_p, unsigned *tickets,
                                                                double *d, int n, int *err, const int *stop)
{
   if (*stop != 0) { return; }
   __shared__ bool s_last;
   const u64 seq = ld_relaxed_gpu(seq_p) + 1;
   const int par = (int)(seq & 1);
   for (int t = 0; t < nranks; t++)
   {
      double *slot = (double *)(peer[t] + vals_off) + ((size_t)par * nranks + rank) * AR_BIG_MAX;
      for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) { slot[k] = d[k]; }
   }
   __syncthreads();
   if (threadIdx.x == 0)
   {
      __threadfence_system();
      s_last = (atomicAdd(tickets, 1u) == gridDim.x - 1);
   }
   __syncthreads();
   if (s_last && threadIdx.x < nranks) { st_release_sys((u64 *)(peer[threadIdx.x] + flag_off) + rank, seq); }
   if (threadIdx.x < nranks) { p2p_spin_ge((const u64 *)(peer[rank] + flag_off) + threadIdx.x, seq, err); }
   __syncthreads();
   const double *mine = (const double *)(peer[rank] + vals_off) + (size_t)par * nranks * AR_BIG_MAX;
   for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x)
   {
      double s = 0.0;
      for (int t = 0; t < nranks; t++) { s += __ldcg(mine + (size_t)t * AR_BIG_MAX + k); }
      d[k] = s;
   }
   __syncthreads();
   if (threadIdx.x == 0)
   {
      // second ticket: the last block to finish advances the sequence and re-arms the tickets
      if (atomicAdd(tickets + 1, 1u) == gridDim.x - 1) { *seq_p = seq; tickets[0] = 0u; tickets[1] = 0u; __threadfence(); }
   }
}

int p2p_allreduce_vec_max(void) { return AR_BIG_MAX; }

int p2p_allreduce_vec(b200_ctx *c, double *d, int n)
{
   if (n > AR_BIG_MAX) { b200_set_error("p2p_allreduce_vec: %d > %d values", n, AR_BIG_MAX); return B200_ERR_ARG; }
   if (n <= 0) { return B200_OK; }
   int grid = (n + 255) / 256;
   if (grid > AR_BIG_GRID) { grid = AR_BIG_GRID; }
   LAUNCH(c, p2p_allreduce_big_kernel, grid, 256, 0, c->d_peer_arena, c->arb_vals_off, c->arb_flag_off, c->rank, c->nranks,
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
