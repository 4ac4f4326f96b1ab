// ctx.cu -- context, device memory, BLAS-1 kernels.
#include "common.cuh"
#include "vecops.cuh"
#include "p2p.cuh"
#include "tail.cuh"
#include <cstdarg>

static thread_local char g_err[1024] = "";

void b200_set_error(const char *fmt, ...)
{
   va_list ap;
   va_start(ap, fmt);
   vsnprintf(g_err, sizeof(g_err), fmt, ap);
   va_end(ap);
}

extern "C" const char *b200_last_error(void) { return g_err; }
extern "C" int b200_version(void) { return 100; }

static int ctx_init_common(b200_ctx *c, int device)
{
   int ndev = 0;
   cudaError_t e = cudaGetDeviceCount(&ndev);
   if (e != cudaSuccess || ndev == 0)
   {
      b200_set_error("no CUDA device available (%s): this library has no CPU fallback",
                     e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
      return B200_ERR_CUDA;
   }
   ARG_CHECK(device >= 0 && device < ndev);
   c->device = device;
   { const char *e = getenv("B200_PROFILE"); c->prof = e ? atoi(e) : 0; }
   CUDA_TRY(cudaSetDevice(device));
   CUDA_TRY(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
   CUDA_TRY(cudaStreamCreateWithFlags(&c->cstream, cudaStreamNonBlocking));
   CUDA_TRY(cudaEventCreateWithFlags(&c->ev_pack, cudaEventDisableTiming));
   CUDA_TRY(cudaEventCreateWithFlags(&c->ev_halo, cudaEventDisableTiming));
   CUDA_TRY(cudaEventCreate(&c->ev_t0));
   CUDA_TRY(cudaEventCreate(&c->ev_t1));
   CUDA_TRY(cudaMalloc((void **)&c->d_partials, sizeof(double) * B200_MAX_RED * B200_MAX_PARTIALS));
   CUDA_TRY(cudaMalloc((void **)&c->d_counter, sizeof(unsigned)));
   CUDA_TRY(cudaMemset(c->d_counter, 0, sizeof(unsigned)));
   CUDA_TRY(cudaMalloc((void **)&c->d_scalars, sizeof(double) * B200_NUM_SCALARS));
   CUDA_TRY(cudaMemset(c->d_scalars, 0, sizeof(double) * B200_NUM_SCALARS));
   CUDA_TRY(cudaMallocHost((void **)&c->h_scalars, sizeof(double) * B200_NUM_SCALARS));
   CUDA_TRY(cudaMalloc((void **)&c->d_stop, sizeof(int)));
   CUDA_TRY(cudaMemset(c->d_stop, 0, sizeof(int)));
   CUDA_TRY(cudaMalloc((void **)&c->d_iter, sizeof(int)));
   CUDA_TRY(cudaMemset(c->d_iter, 0, sizeof(int)));
   CUDA_TRY(cudaHostAlloc(&c->h_ring, 4096, cudaHostAllocMapped));
   memset(c->h_ring, 0, 4096);
   CUDA_TRY(cudaHostGetDevicePointer(&c->d_ring, c->h_ring, 0));
   return B200_OK;
}

// ---------------------------------------------------------------- event profiler
void prof_begin(b200_ctx *c, const char *name)
{
   ProfSample s;
   s.name = name;
   cudaEventCreate(&s.e0); cudaEventCreate(&s.e1);
   cudaEventRecord(s.e0, c->stream);
   c->prof_samples.push_back(s);
}
void prof_end(b200_ctx *c) { cudaEventRecord(c->prof_samples.back().e1, c->stream); }

// text table "kernel  launches  total ms  avg us" of everything launched since the last report
extern "C" int b200_prof_report(b200_ctx *c, char *buf, int len)
{
   ARG_CHECK(c != nullptr && buf != nullptr && len > 0);
   buf[0] = 0;
   CUDA_TRY(cudaStreamSynchronize(c->stream));
   std::vector<std::string> names; std::vector<double> tot; std::vector<long> cnt;
   for (auto &s : c->prof_samples)
   {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, s.e0, s.e1);
      cudaEventDestroy(s.e0); cudaEventDestroy(s.e1);
      size_t k = 0;
      for (; k < names.size(); k++) { if (names[k] == s.name) { break; } }
      if (k == names.size()) { names.push_back(s.name); tot.push_back(0.0); cnt.push_back(0); }
      tot[k] += ms; cnt[k]++;
   }
   c->prof_samples.clear();
   int off = 0;
   double all = 0.0;
   for (double t : tot) { all += t; }
   off += snprintf(buf + off, len - off, "rank %d: %.3f ms in %zu kernel kinds\n", c->rank, all, names.size());
   for (size_t k = 0; k < names.size() && off < len - 200; k++)
   {
      off += snprintf(buf + off, len - off, "  %-60.60s %7ld %10.3f ms %8.1f us\n", names[k].c_str(), cnt[k], tot[k],
                      1e3 * tot[k] / (double)cnt[k]);
   }
   return B200_OK;
}

extern "C" int b200_ctx_create(int device, b200_ctx **ctx)
{
   ARG_CHECK(ctx != nullptr);
   b200_ctx *c = new b200_ctx();
   int s = ctx_init_common(c, device);
   if (s != B200_OK) { delete c; *ctx = nullptr; return s; }
   *ctx = c;
   return B200_OK;
}

extern "C" int b200_nccl_unique_id(char uid[128])
{
   ncclUniqueId id;
   static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
   NCCL_TRY(ncclGetUniqueId(&id));
   memcpy(uid, &id, 128);
   return B200_OK;
}

extern "C" int b200_ctx_create_dist(int device, int rank, int nranks, const char uid[128], b200_ctx **ctx)
{
   ARG_CHECK(ctx != nullptr && nranks >= 1 && rank >= 0 && rank < nranks);
   b200_ctx *c = new b200_ctx();
   int s = ctx_init_common(c, device);
   if (s != B200_OK) { delete c; *ctx = nullptr; return s; }
   c->rank = rank; c->nranks = nranks;
   if (nranks > 1)
   {
      ncclUniqueId id;
      memcpy(&id, uid, 128);
      NCCL_TRY(ncclCommInitRank(&c->comm, nranks, id, rank));
      NCCL_TRY(ncclCommSplit(c->comm, 0, rank, &c->hcomm, nullptr));
      CUDA_TRY(cudaMallocHost((void **)&c->h_err, sizeof(int)));
      *c->h_err = 0;
      s = p2p_bootstrap(c);
      if (s != B200_OK) { *ctx = nullptr; return s; }
   }
   *ctx = c;
   return B200_OK;
}

extern "C" int b200_ctx_destroy(b200_ctx *c)
{
   if (!c) { return B200_OK; }
   cudaSetDevice(c->device);
   cudaDeviceSynchronize();
   p2p_shutdown(c);
   tail_release(c);
   if (c->h_err) { cudaFreeHost(c->h_err); }
   if (c->hcomm) { ncclCommDestroy(c->hcomm); }
   if (c->comm) { ncclCommDestroy(c->comm); }
   cudaFree(c->d_partials); cudaFree(c->d_counter); cudaFree(c->d_scalars);
   cudaFreeHost(c->h_scalars); cudaFree(c->d_stop); cudaFree(c->d_iter);
   if (c->pcg_graph) { cudaGraphExecDestroy(c->pcg_graph); }
   if (c->h_ring) { cudaFreeHost(c->h_ring); }
   if (c->d_flush) { cudaFree(c->d_flush); }
   if (c->h_stage) { cudaFreeHost(c->h_stage); }
   if (c->h_kry) { cudaFreeHost(c->h_kry); }
   for (int k = 0; k < b200_ctx::NUM_WS; k++) { if (c->ws[k]) { cudaFree(c->ws[k]); } }
   cudaEventDestroy(c->ev_pack); cudaEventDestroy(c->ev_halo);
   cudaEventDestroy(c->ev_t0); cudaEventDestroy(c->ev_t1);
   cudaStreamDestroy(c->stream); cudaStreamDestroy(c->cstream);
   delete c;
   return B200_OK;
}

extern "C" int b200_ctx_sync(b200_ctx *c)
{
   ARG_CHECK(c != nullptr);
   CUDA_TRY(cudaStreamSynchronize(c->cstream));
   CUDA_TRY(cudaStreamSynchronize(c->stream));
   return B200_OK;
}

extern "C" int b200_ctx_rank(const b200_ctx *c, int *rank, int *nranks)
{
   ARG_CHECK(c != nullptr);
   if (rank) { *rank = c->rank; }
   if (nranks) { *nranks = c->nranks; }
   return B200_OK;
}

extern "C" long long b200_ctx_launch_count(const b200_ctx *c) { return c ? c->launches : -1; }

extern "C" int b200_timer_start(b200_ctx *c)
{
   ARG_CHECK(c != nullptr);
   CUDA_TRY(cudaEventRecord(c->ev_t0, c->stream));
   return B200_OK;
}

extern "C" int b200_timer_stop(b200_ctx *c, float *ms)
{
   ARG_CHECK(c != nullptr && ms != nullptr);
   CUDA_TRY(cudaEventRecord(c->ev_t1, c->stream));
   CUDA_TRY(cudaEventSynchronize(c->ev_t1));
   CUDA_TRY(cudaEventElapsedTime(ms, c->ev_t0, c->ev_t1));
   return B200_OK;
}

extern "C" int b200_flush_l2(b200_ctx *c)
{
   ARG_CHECK(c != nullptr);
   if (!c->d_flush)
   {
      c->flush_bytes = (size_t)256 << 20;   // 256 MiB > 126 MB L2
      CUDA_TRY(cudaMalloc((void **)&c->d_flush, c->flush_bytes));
   }
   CUDA_TRY(cudaMemsetAsync(c->d_flush, 1, c->flush_bytes, c->stream));
   return B200_OK;
}

// ---------------------------------------------------------------- memory
int ctx_workspace(b200_ctx *c, int slot, size_t n, double **out)
{
   if (n == 0) { n = 1; }
   if (c->ws_n[slot] < n)
   {
      if (c->ws[slot]) { CUDA_TRY(cudaStreamSynchronize(c->stream)); CUDA_TRY(cudaFree(c->ws[slot])); c->ws[slot] = nullptr; c->ws_n[slot] = 0; }
      CUDA_TRY(cudaMalloc((void **)&c->ws[slot], sizeof(double) * n));
      c->ws_n[slot] = n;
   }
   *out = c->ws[slot];
   return B200_OK;
}

// page-lock a host array the caller owns (hypre vectors are plain malloc memory) so that
// the H2D/D2H copies of the host-vector entry points run at full PCIe speed
extern "C" int b200_host_pin(void *host, long long bytes)
{
   ARG_CHECK(host != nullptr && bytes > 0);
   CUDA_TRY(cudaHostRegister(host, (size_t)bytes, cudaHostRegisterDefault));
   return B200_OK;
}
extern "C" int b200_host_unpin(void *host)
{
   ARG_CHECK(host != nullptr);
   CUDA_TRY(cudaHostUnregister(host));
   return B200_OK;
}
extern "C" int b200_dev_alloc(b200_ctx *c, long long n, double **dev)
{
   ARG_CHECK(c != nullptr && dev != nullptr && n >= 0);
   *dev = nullptr;
   CUDA_TRY(cudaMalloc((void **)dev, sizeof(double) * (size_t)(n > 0 ? n : 1)));
   CUDA_TRY(cudaMemsetAsync(*dev, 0, sizeof(double) * (size_t)(n > 0 ? n : 1), c->stream));
   return B200_OK;
}

extern "C" int b200_dev_free(b200_ctx *c, double *dev)
{
   ARG_CHECK(c != nullptr);
   if (dev) { CUDA_TRY(cudaStreamSynchronize(c->stream)); CUDA_TRY(cudaFree(dev)); }
   return B200_OK;
}

extern "C" int b200_h2d(b200_ctx *c, double *dev, const double *host, long long n)
{
   ARG_CHECK(c != nullptr && n >= 0);
   if (n == 0) { return B200_OK; }
   CUDA_TRY(cudaMemcpyAsync(dev, host, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
   CUDA_TRY(cudaStreamSynchronize(c->stream));
   return B200_OK;
}

extern "C" int b200_d2h(b200_ctx *c, double *host, const double *dev, long long n)
{
   ARG_CHECK(c != nullptr && n >= 0);
   if (n == 0) { return B200_OK; }
   CUDA_TRY(cudaMemcpyAsync(host, dev, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
   CUDA_TRY(cudaStreamSynchronize(c->stream));
   return B200_OK;
}

// ---------------------------------------------------------------- BLAS-1
extern "C" int b200_vec_fill(b200_ctx *c, int n, double a, double *y)
{
   ARG_CHECK(c != nullptr && n >= 0);
   return vec_fill(c, n, a, y);
}
extern "C" int b200_vec_copy(b200_ctx *c, int n, const double *x, double *y)
{
   ARG_CHECK(c != nullptr && n >= 0);
   if (n == 0 || x == y) { return B200_OK; }
   if (tail_recording(c))
   {
      TailOp op;
      memset(&op, 0, sizeof(op));
      op.kind = TK_COPY; op.n = n; op.x = x; op.y = y;
      return tail_append(c, op);
   }
   CUDA_TRY(cudaMemcpyAsync(y, x, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, c->stream));
   return B200_OK;
}
extern "C" int b200_vec_scale(b200_ctx *c, int n, double a, double *y)
{
   ARG_CHECK(c != nullptr && n >= 0);
   // hypre_SeqVectorScale: alpha == 1 is a no-op, alpha == 0 fills (src/seq_mv/vector.c:652-689)
   if (a == 1.0) { return B200_OK; }
   if (a == 0.0) { return vec_fill(c, n, 0.0, y); }
   return vec_scale(c, n, a, y);
}
extern "C" int b200_vec_axpy(b200_ctx *c, int n, double a, const double *x, double *y)
{
   ARG_CHECK(c != nullptr && n >= 0);
   return vec_axpy(c, n, a, x, y);
}
extern "C" int b200_vec_dot(b200_ctx *c, int n, const double *x, const double *y, double *result_h)
{
   ARG_CHECK(c != nullptr && n >= 0 && result_h != nullptr);
   B200_TRY(vec_dot(c, n, x, y, c->d_scalars + 0));
   B200_TRY(scalars_allreduce(c, c->d_scalars + 0, 1));
   B200_TRY(scalars_to_host(c, 0, 1));
   *result_h = c->h_scalars[0];
   return B200_OK;
}
