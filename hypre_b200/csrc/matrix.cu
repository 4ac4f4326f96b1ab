// matrix.cu -- upload of CSR blocks into the HBM layouts, ParCSR mirror, halo.
#include "common.cuh"
#include "p2p.cuh"
#include <algorithm>
#include <climits>
#include <cstring>
#include <numeric>

// longest run [out[0], out[1]) of groups whose flag is clear
static void longest_clear_run(const std::vector<unsigned char> &g, int n_groups, int *out)
{
   int best_b = 0, best_e = 0, b = 0;
   for (int k = 0; k <= n_groups; k++)
   {
      if (k == n_groups || g[(size_t)k])
      {
         if (k - b > best_e - best_b) { best_b = b; best_e = k; }
         b = k + 1;
      }
   }
   out[0] = best_b; out[1] = best_e;
}

void dmat_free(DMat *m)
{
   cudaFree(m->rp); cudaFree(m->ci); cudaFree(m->va);
   cudaFree(m->sl_off); cudaFree(m->sl_ci); cudaFree(m->sl_va); cudaFree(m->sl_perm);
   cudaFree(m->c_idx); cudaFree(m->c_base); cudaFree(m->c_code); cudaFree(m->c_dict);
   *m = DMat();
}

// counting-sort transpose; keeps the entries of each transposed row ordered by
// source row (deterministic summation order on the device)
void host_transpose(const HostCsr &a, HostCsr *t)
{
   t->n_rows = a.n_cols; t->n_cols = a.n_rows;
   const size_t nnz = a.j.size();
   t->i.assign((size_t)a.n_cols + 1, 0);
   t->j.resize(nnz); t->a.resize(nnz);
   for (size_t k = 0; k < nnz; k++) { t->i[(size_t)a.j[k] + 1]++; }
   for (int c = 0; c < a.n_cols; c++) { t->i[c + 1] += t->i[c]; }
   std::vector<int> pos(t->i.begin(), t->i.end() - 1);
   for (int r = 0; r < a.n_rows; r++)
   {
      for (int k = a.i[r]; k < a.i[r + 1]; k++)
      {
         int p = pos[a.j[k]]++;
         t->j[p] = r; t->a[p] = a.a[k];
      }
   }
}

// Choose the layout from the row-length statistics (dmat_pack, pack.cu) and upload.
int dmat_upload(b200_ctx *ctx, const HostCsr &h, DMat *m, int allow_sell)
{
   (void)ctx;
   *m = DMat();
   m->n_rows = h.n_rows; m->n_cols = h.n_cols;
   m->nnz = h.n_rows > 0 ? (long long)h.i[h.n_rows] : 0;
   const int n = h.n_rows;
   std::vector<int> zero_rp(1, 0);
   B200_TRY(dev_upload(&m->rp, n > 0 ? h.i.data() : zero_rp.data(), (size_t)n + 1));
   B200_TRY(dev_upload(&m->ci, h.j.data(), (size_t)m->nnz));
   B200_TRY(dev_upload(&m->va, h.a.data(), (size_t)m->nnz));
   if (n == 0) { m->layout = 0; m->lanes = 4; return B200_OK; }

   PackedDMat P;
   dmat_pack(h, allow_sell, &P);
   m->max_row = P.max_row; m->avg_row = P.avg_row;
   m->layout = P.layout; m->lanes = P.lanes;
   if (P.layout == 0) { return B200_OK; }
   m->n_slices = P.n_slices; m->sell_entries = P.padded; m->stored_bytes = P.stored_bytes; m->zpad = P.zpad;
   m->cfmt = P.cfmt; m->n_dict = (int)P.dict.size();
   B200_TRY(dev_upload(&m->sl_off, P.sl_off.data(), P.sl_off.size()));
   B200_TRY(dev_upload(&m->sl_ci, P.sci.data(), P.sci.size()));
   B200_TRY(dev_upload(&m->sl_va, P.sva.data(), P.sva.size()));
   B200_TRY(dev_upload(&m->sl_perm, P.perm.data(), P.perm.size()));
   if (P.cfmt != CFMT_PAT8) { B200_TRY(dev_upload(&m->c_idx, P.idx.data(), P.idx.size())); }
   B200_TRY(dev_upload(&m->c_base, P.base.data(), P.base.size()));
   unsigned char *d_code = nullptr;
   if (P.cfmt == CFMT_PAT8)
   {
      B200_TRY(dev_upload(&d_code, P.pat.data(), P.pat.size()));
      unsigned *d_lanes = nullptr;
      B200_TRY(dev_upload(&d_lanes, P.pat_lanes.data(), P.pat_lanes.size()));
      m->c_idx = reinterpret_cast<unsigned short *>(d_lanes);      // lane words live in the c_idx slot
      m->n_dict = (int)(P.dict.size() / 2);
      m->pat_stride = P.pat_stride;
   }
   else { B200_TRY(dev_upload(&d_code, P.code.data(), P.code.size())); }
   m->c_code = d_code;
   B200_TRY(dev_upload(&m->c_dict, P.dict.data(), P.dict.size()));
   m->h_pos.swap(P.h_pos);
   return B200_OK;
}

static void fill_host_csr(HostCsr *h, int n_rows, int n_cols, const int *i, const int *j, const double *a)
{
   h->n_rows = n_rows; h->n_cols = n_cols;
   if (n_rows == 0 || i == nullptr) { h->i.assign((size_t)n_rows + 1, 0); h->j.clear(); h->a.clear(); return; }
   h->i.assign(i, i + n_rows + 1);
   const size_t nnz = (size_t)i[n_rows];
   h->j.assign(j, j + nnz);
   h->a.assign(a, a + nnz);
}

extern "C" int b200_parcsr_create(b200_ctx *ctx, int n_rows, int n_cols_diag,
                                  const int *diag_i, const int *diag_j, const double *diag_a,
                                  int n_cols_offd,
                                  const int *offd_i, const int *offd_j, const double *offd_a,
                                  int num_sends, const int *send_procs, const int *send_map_starts,
                                  const int *send_map_elmts,
                                  int num_recvs, const int *recv_procs, const int *recv_vec_starts,
                                  b200_parcsr **out)
{
   ARG_CHECK(ctx != nullptr && out != nullptr && n_rows >= 0 && n_cols_diag >= 0 && n_cols_offd >= 0);
   ARG_CHECK(n_rows == 0 || diag_i != nullptr);
   ARG_CHECK(num_sends >= 0 && num_recvs >= 0);
   CUDA_TRY(cudaSetDevice(ctx->device));
   b200_parcsr *A = new b200_parcsr();
   A->ctx = ctx; A->n_rows = n_rows; A->n_cols_diag = n_cols_diag; A->n_cols_offd = n_cols_offd;
   fill_host_csr(&A->h_diag, n_rows, n_cols_diag, diag_i, diag_j, diag_a);
   fill_host_csr(&A->h_offd, n_rows, n_cols_offd, n_cols_offd > 0 ? offd_i : nullptr, offd_j, offd_a);
   int s = dmat_upload(ctx, A->h_diag, &A->diag, 1);
   if (s == B200_OK) { s = dmat_upload(ctx, A->h_offd, &A->offd, 0); }
   if (s != B200_OK) { b200_parcsr_destroy(A); return s; }

   A->num_sends = num_sends; A->num_recvs = num_recvs;
   if (num_sends > 0)
   {
      A->send_procs.assign(send_procs, send_procs + num_sends);
      A->send_map_starts.assign(send_map_starts, send_map_starts + num_sends + 1);
      A->n_send_elmts = send_map_starts[num_sends];
      s = dev_upload(&A->d_send_map, send_map_elmts, (size_t)A->n_send_elmts);
      if (s == B200_OK && A->n_send_elmts > 0)
      {
         if (cudaMalloc((void **)&A->d_send_buf, sizeof(double) * A->n_send_elmts) != cudaSuccess ||
             cudaMalloc((void **)&A->d_recv_buf, sizeof(double) * A->n_send_elmts) != cudaSuccess) { s = B200_ERR_CUDA; }
      }
      // deterministic job-2 unpack plan: positions of the receive buffer grouped by target row
      if (s == B200_OK && A->n_send_elmts > 0)
      {
         std::vector<int> order(A->n_send_elmts);
         std::iota(order.begin(), order.end(), 0);
         std::stable_sort(order.begin(), order.end(),
                          [&](int a, int b) { return send_map_elmts[a] < send_map_elmts[b]; });
         std::vector<int> rows, rp(1, 0);
         for (int k = 0; k < A->n_send_elmts; k++)
         {
            int r = send_map_elmts[order[k]];
            if (rows.empty() || rows.back() != r) { if (!rows.empty()) { rp.push_back(k); } rows.push_back(r); }
         }
         rp.push_back(A->n_send_elmts);
         A->n_unpack_rows = (int)rows.size();
         s = dev_upload(&A->d_unpack_rows, rows.data(), rows.size());
         if (s == B200_OK) { s = dev_upload(&A->d_unpack_rp, rp.data(), rp.size()); }
         if (s == B200_OK) { s = dev_upload(&A->d_unpack_pos, order.data(), order.size()); }
         if (s == B200_OK)
         {
            std::vector<int> dense((size_t)n_cols_diag + 1, 0);
            for (int k = 0; k < A->n_send_elmts; k++) { dense[(size_t)send_map_elmts[k] + 1]++; }
            for (int r = 0; r < n_cols_diag; r++) { dense[(size_t)r + 1] += dense[(size_t)r]; }
            s = dev_upload(&A->d_unpack_rp_dense, dense.data(), dense.size());
         }
      }
   }
   if (s == B200_OK && num_recvs > 0)
   {
      A->recv_procs.assign(recv_procs, recv_procs + num_recvs);
      A->recv_vec_starts.assign(recv_vec_starts, recv_vec_starts + num_recvs + 1);
   }
   if (s == B200_OK && n_cols_offd > 0)
   {
      if (cudaMalloc((void **)&A->d_ghost, sizeof(double) * n_cols_offd) != cudaSuccess ||
          cudaMalloc((void **)&A->d_ghost_out, sizeof(double) * n_cols_offd) != cudaSuccess) { s = B200_ERR_CUDA; }
      std::vector<int> bnd;
      for (int r = 0; r < n_rows; r++) { if (A->h_offd.i[r + 1] > A->h_offd.i[r]) { bnd.push_back(r); } }
      A->n_bnd = (int)bnd.size();
      if (s == B200_OK) { s = dev_upload(&A->d_bnd_rows, bnd.data(), bnd.size()); }
   }
   if (s != B200_OK)
   {
      if (s == B200_ERR_CUDA) { b200_set_error("device allocation failed in b200_parcsr_create"); }
      b200_parcsr_destroy(A);
      return s;
   }
   if (ctx->use_p2p && (num_sends + num_recvs) > 0)
   {
      // 32-row STORAGE groups of the diag block that contain a row with offd entries (layout 2 stores
      // the rows of a window in sorted order: use the storage position)
      std::vector<unsigned char> gf((size_t)(n_rows + 31) / 32 + 1, 0);
      if (n_cols_offd > 0)
      {
         for (int r = 0; r < n_rows; r++)
         {
            if (A->h_offd.i[r + 1] > A->h_offd.i[r])
            {
               const int pos = A->diag.layout == 2 ? A->diag.h_pos[(size_t)r] : r;
               gf[(size_t)pos >> 5] = 1;
            }
         }
      }
      longest_clear_run(gf, (n_rows + 31) / 32, A->int_fwd);
      s = dev_upload(&A->d_grp_fwd, gf.data(), gf.size());
      if (s != B200_OK) { b200_parcsr_destroy(A); return s; }
      if (A->d_unpack_rp_dense == nullptr)
      {
         // this rank receives no contributions for A^T x: the fused kernel still reads a per-row list
         std::vector<int> dense((size_t)n_cols_diag + 1, 0);
         s = dev_upload(&A->d_unpack_rp_dense, dense.data(), dense.size());
         if (s != B200_OK) { b200_parcsr_destroy(A); return s; }
      }
      // forward: I write x[send_map] into the ghost vectors of send_procs, recv_procs write into mine;
      // reverse (A^T x): roles swapped, segments = the same starts
      const int *sms = num_sends > 0 ? A->send_map_starts.data() : nullptr;
      const int *rvs = num_recvs > 0 ? A->recv_vec_starts.data() : nullptr;
      s = p2p_chan_create(ctx, num_sends, A->send_procs.data(), sms, num_recvs, A->recv_procs.data(), rvs, &A->ch_fwd);
      if (s == B200_OK)
      {
         s = p2p_chan_create(ctx, num_recvs, A->recv_procs.data(), rvs, num_sends, A->send_procs.data(), sms, &A->ch_rev);
      }
      if (s != B200_OK) { b200_parcsr_destroy(A); return s; }
   }
   *out = A;
   return B200_OK;
}

extern "C" int b200_parcsr_destroy(b200_parcsr *A)
{
   if (!A) { return B200_OK; }
   if (A->ctx) { A->ctx->generation++; }
   if (A->ctx) { cudaSetDevice(A->ctx->device); cudaStreamSynchronize(A->ctx->stream); cudaStreamSynchronize(A->ctx->cstream); }
   p2p_chan_destroy(A->ch_fwd); p2p_chan_destroy(A->ch_rev);
   dmat_free(&A->diag); dmat_free(&A->offd);
   dmat_free(&A->diagT); dmat_free(&A->offdT); dmat_free(&A->lower);
   cudaFree(A->d_diagonal);
   cudaFree(A->d_send_map); cudaFree(A->d_send_buf); cudaFree(A->d_recv_buf);
   cudaFree(A->d_ghost); cudaFree(A->d_ghost_out); cudaFree(A->d_bnd_rows);
   cudaFree(A->d_unpack_rp); cudaFree(A->d_unpack_pos); cudaFree(A->d_unpack_rows); cudaFree(A->d_unpack_rp_dense); cudaFree(A->d_grp_fwd); cudaFree(A->d_grp_rev);
   delete A;
   return B200_OK;
}

extern "C" int b200_parcsr_info(const b200_parcsr *A, int *n_rows, int *n_cols, long long *nnz,
                                int *layout, int *lanes_per_row)
{
   ARG_CHECK(A != nullptr);
   if (n_rows) { *n_rows = A->n_rows; }
   if (n_cols) { *n_cols = A->n_cols_diag; }
   if (nnz)    { *nnz = A->diag.nnz + A->offd.nnz; }
   if (layout) { *layout = A->diag.layout; }
   if (lanes_per_row) { *lanes_per_row = A->diag.lanes; }
   return B200_OK;
}

extern "C" int b200_parcsr_storage(const b200_parcsr *A, int *value_format, long long *stored_bytes,
                                   long long *csr_bytes, int *n_dictionary)
{
   ARG_CHECK(A != nullptr);
   const DMat &D = A->diag;
   const long long csr = 12 * D.nnz + 4 * ((long long)D.n_rows + 1);
   if (value_format) { *value_format = D.cfmt; }
   if (stored_bytes) { *stored_bytes = D.layout >= 1 ? D.stored_bytes : csr; }
   if (csr_bytes)    { *csr_bytes = csr; }
   if (n_dictionary) { *n_dictionary = D.n_dict; }
   return B200_OK;
}

// explicit A^T blocks for restriction (R = P^T, src/parcsr_ls/par_amg_setup.c:819)
int parcsr_ensure_T(b200_parcsr *A)
{
   if (A->hasT) { return B200_OK; }
   HostCsr t;
   host_transpose(A->h_diag, &t);
   B200_TRY(dmat_upload(A->ctx, t, &A->diagT, 1));
   if (A->n_cols_offd > 0)
   {
      HostCsr to;
      host_transpose(A->h_offd, &to);
      B200_TRY(dmat_upload(A->ctx, to, &A->offdT, 0));
   }
   if (A->ctx->use_p2p && (A->num_sends + A->num_recvs) > 0)
   {
      // storage groups of diag^T whose rows receive contributions from other ranks (A^T x, job 2)
      const int nr = A->n_cols_diag;
      std::vector<unsigned char> gr((size_t)(nr + 31) / 32 + 1, 0);
      std::vector<int> sme((size_t)A->n_send_elmts);
      if (A->n_send_elmts > 0)
      {
         CUDA_TRY(cudaMemcpy(sme.data(), A->d_send_map, sizeof(int) * sme.size(), cudaMemcpyDeviceToHost));
      }
      for (int k = 0; k < A->n_send_elmts; k++)
      {
         const int pos = A->diagT.layout == 2 ? A->diagT.h_pos[(size_t)sme[(size_t)k]] : sme[(size_t)k];
         gr[(size_t)pos >> 5] = 1;
      }
      longest_clear_run(gr, (nr + 31) / 32, A->int_rev);
      cudaFree(A->d_grp_rev); A->d_grp_rev = nullptr;
      B200_TRY(dev_upload(&A->d_grp_rev, gr.data(), gr.size()));
   }
   A->hasT = true;
   return B200_OK;
}

// Owner blocks of a gathered (replicated) level: the strict lower triangle used by two-stage Gauss-Seidel is then
// that of the owners' diagonal blocks, as in the reference's distributed sweep (par_relax.c:1505-1588 works on the
// rank's diag block only; couplings to other ranks' rows are treated Jacobi-like).  rows_per_block[b] = rows of block b.
extern "C" int b200_parcsr_set_gs_blocks(b200_parcsr *A, int n_blocks, const int *rows_per_block)
{
   ARG_CHECK(A != nullptr && n_blocks >= 1 && rows_per_block != nullptr && !A->hasL);
   A->gs_block_of.assign((size_t)A->n_rows, 0);
   int r = 0;
   for (int b = 0; b < n_blocks; b++) { for (int k = 0; k < rows_per_block[b] && r < A->n_rows; k++) { A->gs_block_of[(size_t)r++] = b; } }
   ARG_CHECK(r == A->n_rows);
   return B200_OK;
}

// strict lower triangle of the diag block (local column < row) and the diagonal
// (first entry of each diag row by hypre's convention; we search for it so that a
// missing diagonal is detected) for the two-stage Gauss-Seidel smoother
// (src/parcsr_ls/par_relax.c:1538-1581).
int parcsr_ensure_L(b200_parcsr *A)
{
   if (A->hasL) { return B200_OK; }
   const HostCsr &h = A->h_diag;
   HostCsr L;
   L.n_rows = h.n_rows; L.n_cols = h.n_cols;
   L.i.assign((size_t)h.n_rows + 1, 0);
   std::vector<double> d((size_t)h.n_rows, 0.0);
   A->zero_diagonal = false;
   for (int r = 0; r < h.n_rows; r++)
   {
      // the reference divides by A_diag_data[A_diag_i[i]] (first entry of the row)
      double dv = (h.i[r + 1] > h.i[r]) ? h.a[h.i[r]] : 0.0;
      d[r] = dv;
      if (dv == 0.0) { A->zero_diagonal = true; }
      for (int k = h.i[r]; k < h.i[r + 1]; k++)
      {
         const bool same_block = A->gs_block_of.empty() || (h.j[k] < h.n_rows && A->gs_block_of[(size_t)h.j[k]] == A->gs_block_of[(size_t)r]);
         if (h.j[k] < r && same_block) { L.j.push_back(h.j[k]); L.a.push_back(h.a[k]); }
      }
      L.i[r + 1] = (int)L.j.size();
   }
   B200_TRY(dmat_upload(A->ctx, L, &A->lower, 1));
   B200_TRY(dev_upload(&A->d_diagonal, d.data(), d.size()));
   A->hasL = true;
   return B200_OK;
}

// ---------------------------------------------------------------- halo (job 1)
// pack: x_buf[i] = x_local[send_map_elmts[i]]  (src/parcsr_mv/par_csr_matvec.c:170-175)
__global__ void halo_pack_kernel(int n, const int *__restrict__ map, const double *__restrict__ x,
                                 double *__restrict__ buf)
{
   int i = blockIdx.x * blockDim.x + threadIdx.x;
   if (i < n) { buf[i] = x[map[i]]; }
}

// Replaces hypre_ParCSRCommHandleCreate_v2(job 1) = Irecv/Isend per neighbour
// (src/parcsr_mv/par_csr_communication.c:483-504) by one NCCL group on the comm
// stream; the compute stream keeps running the interior rows meanwhile.
int halo_start(b200_parcsr *A, const double *x)
{
   b200_ctx *c = A->ctx;
   if (c->nranks <= 1 || (A->num_sends + A->num_recvs) == 0) { return B200_OK; }
   if (A->n_send_elmts > 0)
   {
      int grid = (A->n_send_elmts + 255) / 256;
      LAUNCH(c, halo_pack_kernel, grid, 256, 0, A->n_send_elmts, A->d_send_map, x, A->d_send_buf);
      CHECK_LAUNCH();
   }
   CUDA_TRY(cudaEventRecord(c->ev_pack, c->stream));
   CUDA_TRY(cudaStreamWaitEvent(c->cstream, c->ev_pack, 0));
   NCCL_TRY(ncclGroupStart());
   for (int i = 0; i < A->num_recvs; i++)
   {
      int cnt = A->recv_vec_starts[i + 1] - A->recv_vec_starts[i];
      NCCL_TRY(ncclRecv(A->d_ghost + A->recv_vec_starts[i], (size_t)cnt, ncclDouble, A->recv_procs[i], c->hcomm, c->cstream));
   }
   for (int i = 0; i < A->num_sends; i++)
   {
      int cnt = A->send_map_starts[i + 1] - A->send_map_starts[i];
      NCCL_TRY(ncclSend(A->d_send_buf + A->send_map_starts[i], (size_t)cnt, ncclDouble, A->send_procs[i], c->hcomm, c->cstream));
   }
   NCCL_TRY(ncclGroupEnd());
   CUDA_TRY(cudaEventRecord(c->ev_halo, c->cstream));
   return B200_OK;
}

// hypre_ParCSRCommHandleDestroy = MPI_Waitall (par_csr_communication.c:653-663):
// here a stream dependency, the host never blocks.
int halo_finish(b200_parcsr *A)
{
   b200_ctx *c = A->ctx;
   if (c->nranks <= 1 || (A->num_sends + A->num_recvs) == 0) { return B200_OK; }
   CUDA_TRY(cudaStreamWaitEvent(c->stream, c->ev_halo, 0));
   return B200_OK;
}

// ---------------------------------------------------------------- halo (job 2)
// y_local[send_map_elmts[i]] += y_buf[i]  (src/parcsr_mv/par_csr_matvec.c:491-496),
// grouped by target row so the summation order is fixed.
__global__ void halo_unpack_add_kernel(int n_rows, const int *__restrict__ rows, const int *__restrict__ rp,
                                       const int *__restrict__ pos, const double *__restrict__ buf,
                                       double alpha, double *__restrict__ y)
{
   int i = blockIdx.x * blockDim.x + threadIdx.x;
   if (i < n_rows)
   {
      double t = 0.0;
      for (int k = rp[i]; k < rp[i + 1]; k++) { t += buf[pos[k]]; }
      y[rows[i]] += alpha * t;
   }
}
// send the ghost-column contributions back to their owners (reverse of job 1)
int halo_reverse_start(b200_parcsr *A)
{
   b200_ctx *c = A->ctx;
   if (c->nranks <= 1 || (A->num_sends + A->num_recvs) == 0) { return B200_OK; }
   CUDA_TRY(cudaEventRecord(c->ev_pack, c->stream));
   CUDA_TRY(cudaStreamWaitEvent(c->cstream, c->ev_pack, 0));
   NCCL_TRY(ncclGroupStart());
   for (int i = 0; i < A->num_sends; i++)
   {
      int cnt = A->send_map_starts[i + 1] - A->send_map_starts[i];
      NCCL_TRY(ncclRecv(A->d_recv_buf + A->send_map_starts[i], (size_t)cnt, ncclDouble, A->send_procs[i], c->hcomm, c->cstream));
   }
   for (int i = 0; i < A->num_recvs; i++)
   {
      int cnt = A->recv_vec_starts[i + 1] - A->recv_vec_starts[i];
      NCCL_TRY(ncclSend(A->d_ghost_out + A->recv_vec_starts[i], (size_t)cnt, ncclDouble, A->recv_procs[i], c->hcomm, c->cstream));
   }
   NCCL_TRY(ncclGroupEnd());
   CUDA_TRY(cudaEventRecord(c->ev_halo, c->cstream));
   return B200_OK;
}

int halo_reverse_finish_add(b200_parcsr *A, double alpha, double *y)
{
   b200_ctx *c = A->ctx;
   if (c->nranks <= 1 || (A->num_sends + A->num_recvs) == 0) { return B200_OK; }
   CUDA_TRY(cudaStreamWaitEvent(c->stream, c->ev_halo, 0));
   if (A->n_unpack_rows > 0)
   {
      int grid = (A->n_unpack_rows + 255) / 256;
      LAUNCH(c, halo_unpack_add_kernel, grid, 256, 0, A->n_unpack_rows, A->d_unpack_rows, A->d_unpack_rp,
             A->d_unpack_pos, A->d_recv_buf, alpha, y);
      CHECK_LAUNCH();
   }
   return B200_OK;
}
