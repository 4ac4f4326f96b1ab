// p2p.cuh -- neighbour exchange and small all-reduce over NVLink peer memory.
//
// Replaces the MPI_Irecv/Isend/Waitall of hypre_ParCSRCommHandleCreate_v2/Destroy
// (src/parcsr_mv/par_csr_communication.c:483-504,653-663) and the MPI_Allreduce of
// hypre_ParVectorInnerProd (src/parcsr_mv/par_vector.c:526) by plain stores into the
// receiver's HBM: every rank owns an "arena" (one cudaMalloc block, CUDA-IPC mapped by all
// peers); a sender's pack kernel gathers x[send_map] and writes it straight into the
// receiver's ghost vector through NVLink, then publishes a sequence number with a
// system-scope release store; the receiver's boundary-row kernel spins on that number with
// acquire loads before it touches the ghosts.  No library call, no extra stream, no host
// involvement per exchange -- one exchange costs one small kernel on the sender.
//
// Flow control: ghost buffers are double buffered (parity of the sequence number).  A sender overwrites the buffer
// of exchange s with exchange s + 2; between the two it completed exchange s + 1, for which it had to RECEIVE the
// peer's data of s + 1 -- which the peer sends at the start of its own kernel s + 1, i.e. after its kernel s (the
// reader of buffer s) ended.  So between peers that exchange in both directions (every neighbour pair of a
// structurally symmetric operator) consumption is implied and nothing is acknowledged; a receiver that does not
// send to its sender acknowledges with its own sequence number and that sender waits for it (per pair).  Sequence counters live in device memory and are advanced
// by the last block of the kernels themselves, so a launch sequence is replayable (CUDA graphs).
// Every spin loop is bounded (P2P_SPIN_LIMIT clock ticks); on expiry the kernel sets the
// context's device error word and continues, so a protocol bug can never hang the GPU.
#pragma once
#include "common.cuh"

typedef unsigned long long u64;

constexpr long long P2P_SPIN_LIMIT = 60000000000LL;   // ~30 s at 1.9 GHz

// Flag-in-data transport: every value travels as 16 bytes -- {low word, sequence} {high word, sequence} -- written
// with two 8-byte (single-copy atomic) stores; the receiver spins on the element it needs until both halves carry
// the sequence number of this exchange.  No fence, no separate flag store, no flag wait: the latency of an exchange
// is ONE NVLink store latency (the put -> fence.sys -> flag -> poll chain of round 1 cost ~10 us per operator, which
// is what bounded strong scaling on the small levels).  The wire carries twice the bytes -- it was 0.2 % busy.
struct __align__(16) LLVal { u64 lo, hi; };

// One direction of a neighbour exchange.  "out" = peers I write to, "in" = peers that write to me.
struct P2PChan {
   int      n_out, n_in;
   int      n_out_total;          // packed elements I send
   const int *out_starts;         // [n_out+1] segment starts in packed order          (device)
   LLVal  **out_dst[2];           // [n_out]  remote data pointers, one per parity     (device array of peer ptrs)
   u64     *out_ack;              // [n_out]  local: last sequence consumed by that receiver (written remotely)
   const int *out_wait_ack;       // [n_out]  1: that receiver does not send to me, so nothing else tells me that it
                                  //          consumed exchange s before I overwrite its buffer with s + 2 -> wait for its ack
   LLVal   *in_buf[2];            // local receive buffers (in my arena)
   u64    **in_ack;               // [n_in]   remote ack word at that sender            (device array of peer ptrs)
   const int *in_send_ack;        // [n_in]   1: that sender waits for my acknowledgement (I do not send to it)
   int      n_ack_in;             // senders that want acknowledgements (0 for symmetric neighbourhoods: no ack traffic)
   u64     *seq_put;              // local counters advanced by the kernels
   u64     *seq_wait;
   unsigned *ticket;              // [2] block tickets (put kernel, consumer kernel)
   int     *err;                  // context error word
};

__device__ __forceinline__ void ll_store(LLVal *dst, double v, unsigned flag)
{
   const u64 b = (u64)__double_as_longlong(v);
   const u64 f = (u64)flag << 32;
   asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" :: "l"(dst), "l"((b & 0xFFFFFFFFull) | f), "l"((b >> 32) | f) : "memory");
}
// spin (bounded) until the element carries `flag`
__device__ __forceinline__ double ll_read(const LLVal *src, unsigned flag, int *err)
{
   u64 lo, hi;
   asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(lo), "=l"(hi) : "l"(src) : "memory");
   if ((unsigned)(lo >> 32) != flag || (unsigned)(hi >> 32) != flag)
   {
      const long long t0 = clock64();
      do
      {
         asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(lo), "=l"(hi) : "l"(src) : "memory");
         if (clock64() - t0 > P2P_SPIN_LIMIT) { atomicExch(err, 1); break; }
      } while ((unsigned)(lo >> 32) != flag || (unsigned)(hi >> 32) != flag);
   }
   return __longlong_as_double((long long)((lo & 0xFFFFFFFFull) | (hi << 32)));
}

__device__ __forceinline__ u64 ld_acquire_sys(const u64 *p)
{
   u64 v;
   asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
   return v;
}
__device__ __forceinline__ void st_release_sys(u64 *p, u64 v)
{
   asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(u64 *p, u64 v)
{
   asm volatile("st.relaxed.sys.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ u64 ld_relaxed_sys(const u64 *p)
{
   u64 v;
   asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
   return v;
}
__device__ __forceinline__ u64 ld_relaxed_gpu(const u64 *p)
{
   u64 v;
   asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
   return v;
}

// spin until *p >= want (bounded)
__device__ __forceinline__ void p2p_spin_ge(const u64 *p, u64 want, int *err)
{
   if (ld_acquire_sys(p) >= want) { return; }
   const long long t0 = clock64();
   while (ld_relaxed_sys(p) < want)
   {
      if (clock64() - t0 > P2P_SPIN_LIMIT) { atomicExch(err, 1); break; }
   }
   (void)ld_acquire_sys(p);      // order the data reads after the observed sequence number
}

// Consumer side, end of kernel: the last block acknowledges to every sender and advances the
// local wait counter.  Call with all threads of the block after the last read of the buffer.
__device__ __forceinline__ void p2p_consumer_done(const P2PChan &ch)
{
   __shared__ bool s_last_c;
   __syncthreads();
   if (threadIdx.x == 0)
   {
      __threadfence();
      s_last_c = (atomicAdd(ch.ticket + 1, 1u) == gridDim.x - 1);
   }
   __syncthreads();
   if (s_last_c)
   {
      // every block finished reading the buffer before it took its ticket (barrier + fence above), so
      // the acknowledgement needs no further fence: a plain system-scope store.  Only senders that cannot infer the
      // consumption from my own next put (I do not send to them) are acknowledged.
      const u64 seq = ld_relaxed_gpu(ch.seq_wait) + 1;
      if (ch.n_ack_in > 0)
      {
         for (int j = threadIdx.x; j < ch.n_in; j += blockDim.x) { if (ch.in_send_ack[j]) { st_relaxed_sys(ch.in_ack[j], seq); } }
      }
      if (threadIdx.x == 0) { *ch.seq_wait = seq; ch.ticket[1] = 0u; }
   }
}

struct b200_p2p_chan;   // host-side owner of a P2PChan (p2p.cu)

// host API (p2p.cu)
int  p2p_bootstrap(b200_ctx *c);                     // arena + IPC exchange (after the NCCL communicator exists)
void p2p_shutdown(b200_ctx *c);
// collective over the participating neighbours: builds one direction of an exchange.
//   out_procs/out_starts: peers I send to and the segment boundaries of my packed data
//   in_procs/in_starts:   peers I receive from and where their data lands in my buffer
int  p2p_chan_create(b200_ctx *c, int n_out, const int *out_procs, const int *out_starts,
                     int n_in, const int *in_procs, const int *in_starts, b200_p2p_chan **out);
void p2p_chan_destroy(b200_p2p_chan *ch);
const P2PChan &p2p_chan_view(const b200_p2p_chan *ch);
// The put / wait / acknowledge steps themselves are device code inside the row kernels (spmv.cuh:
// halo_put_phase, halo_warp_wait, p2p_consumer_done): one launch per operator.
// d[0..cnt) <- sum over ranks, deterministic rank order, result identical on all ranks
int  p2p_allreduce(b200_ctx *c, double *d, int cnt);
// vector version (n <= p2p_allreduce_vec_max()), several blocks
int  p2p_allreduce_vec(b200_ctx *c, double *d, int n);
int  p2p_allreduce_vec_max(void);
// stream-ordered rendezvous of all ranks without a time limit (NCCL): called once at the start of a
// solve so that ranks that are still uploading do not trip the bounded spins of the peers
int  ctx_rendezvous(b200_ctx *c);
