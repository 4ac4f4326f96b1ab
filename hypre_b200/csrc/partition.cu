// partition.cu -- host-side 1-D row partitioning of a replicated CSR matrix into the
// ParCSR form (diag + offd + col_map_offd + halo plan) of one rank.
//
// Every rank holds the same global CSR (the level operators of a hierarchy that the
// reference's host setup produced once); rank r keeps rows [row_starts[r], row_starts[r+1])
// and splits them exactly the way the reference does:
//   * diag/offd split and compressed, sorted col_map_offd:
//       GenerateDiagAndOffd            (src/parcsr_mv/par_csr_matrix.c:2088-2230)
//       hypre_CSRMatrixToParCSRMatrix  (src/parcsr_mv/par_csr_matrix.c:1812-2080)
//   * neighbour lists / pack map of hypre_ParCSRCommPkg:
//       hypre_MatvecCommPkgCreate      (src/parcsr_mv/par_csr_communication.c:1163-1200)
//       (recv blocks ordered by owner rank, send_map_elmts in the order of the
//        requester's col_map_offd, i.e. ascending global column)
// No MPI is needed because each rank can see every other rank's rows; the plan it derives
// for itself is the one the reference would have negotiated with Allgather/exchange.
// Pure host code: usable (and tested) without a GPU.
#include "common.cuh"
#include <algorithm>

struct b200_rowpart {
   int rank = 0, nranks = 1;
   int n_rows = 0, n_cols_diag = 0, n_cols_offd = 0;
   int first_row = 0, first_col = 0;
   std::vector<int>    diag_i, diag_j, offd_i, offd_j, col_map_offd;
   std::vector<double> diag_a, offd_a;
   std::vector<int>    send_procs, send_map_starts, send_map_elmts, recv_procs, recv_vec_starts;
};

static inline int owner_of(const int *starts, int nranks, int g)
{
   // starts[] is non-decreasing; owner r satisfies starts[r] <= g < starts[r+1]
   int r = (int)(std::upper_bound(starts, starts + nranks + 1, g) - starts) - 1;
   return r < 0 ? 0 : (r >= nranks ? nranks - 1 : r);
}

extern "C" int b200_rowpart_create(int rank, int nranks, int n_rows, int n_cols,
                                   const int *gi, const int *gj, const double *ga,
                                   const int *row_starts, const int *col_starts, b200_rowpart **out)
{
   ARG_CHECK(out != nullptr && nranks >= 1 && rank >= 0 && rank < nranks && n_rows >= 0 && n_cols >= 0);
   ARG_CHECK(row_starts != nullptr && col_starts != nullptr);
   ARG_CHECK(n_rows == 0 || (gi != nullptr));
   ARG_CHECK(row_starts[0] == 0 && row_starts[nranks] == n_rows && col_starts[0] == 0 && col_starts[nranks] == n_cols);
   for (int r = 0; r < nranks; r++) { ARG_CHECK(row_starts[r] <= row_starts[r + 1] && col_starts[r] <= col_starts[r + 1]); }

   b200_rowpart *p = new b200_rowpart();
   p->rank = rank; p->nranks = nranks;
   const int r0 = row_starts[rank], r1 = row_starts[rank + 1];
   const int c0 = col_starts[rank], c1 = col_starts[rank + 1];
   p->first_row = r0; p->first_col = c0;
   p->n_rows = r1 - r0; p->n_cols_diag = c1 - c0;

   // ---- ghost columns of my rows (sorted, unique) -> col_map_offd
   std::vector<int> ghosts;
   for (int i = r0; i < r1; i++)
   {
      for (int k = gi[i]; k < gi[i + 1]; k++) { const int c = gj[k]; if (c < c0 || c >= c1) { ghosts.push_back(c); } }
   }
   std::sort(ghosts.begin(), ghosts.end());
   ghosts.erase(std::unique(ghosts.begin(), ghosts.end()), ghosts.end());
   p->col_map_offd = ghosts;
   p->n_cols_offd = (int)ghosts.size();

   // ---- split, keeping the storage order of each row inside its block
   p->diag_i.assign((size_t)p->n_rows + 1, 0);
   p->offd_i.assign((size_t)p->n_rows + 1, 0);
   for (int i = r0; i < r1; i++)
   {
      for (int k = gi[i]; k < gi[i + 1]; k++)
      {
         const int c = gj[k];
         if (c < c0 || c >= c1)
         {
            p->offd_j.push_back((int)(std::lower_bound(ghosts.begin(), ghosts.end(), c) - ghosts.begin()));
            p->offd_a.push_back(ga[k]);
         }
         else { p->diag_j.push_back(c - c0); p->diag_a.push_back(ga[k]); }
      }
      p->diag_i[i - r0 + 1] = (int)p->diag_j.size();
      p->offd_i[i - r0 + 1] = (int)p->offd_j.size();
   }

   // ---- receives: ghost columns grouped by owner (col_map_offd is sorted, owners are contiguous)
   p->recv_vec_starts.push_back(0);
   for (int g = 0; g < p->n_cols_offd; g++)
   {
      const int q = owner_of(col_starts, nranks, ghosts[g]);
      if (p->recv_procs.empty() || p->recv_procs.back() != q)
      {
         if (!p->recv_procs.empty()) { p->recv_vec_starts.push_back(g); }
         p->recv_procs.push_back(q);
      }
   }
   if (p->n_cols_offd > 0) { p->recv_vec_starts.push_back(p->n_cols_offd); }

   // ---- sends: for every other rank q, the columns in [c0, c1) its rows reference (sorted)
   p->send_map_starts.push_back(0);
   if (c1 > c0)
   {
      std::vector<char> mark((size_t)(c1 - c0));
      for (int q = 0; q < nranks; q++)
      {
         if (q == rank) { continue; }
         std::fill(mark.begin(), mark.end(), 0);
         bool any = false;
         for (int i = row_starts[q]; i < row_starts[q + 1]; i++)
         {
            for (int k = gi[i]; k < gi[i + 1]; k++)
            {
               const int c = gj[k];
               if (c >= c0 && c < c1) { mark[(size_t)(c - c0)] = 1; any = true; }
            }
         }
         if (!any) { continue; }
         p->send_procs.push_back(q);
         for (int c = 0; c < c1 - c0; c++) { if (mark[(size_t)c]) { p->send_map_elmts.push_back(c); } }
         p->send_map_starts.push_back((int)p->send_map_elmts.size());
      }
   }
   *out = p;
   return B200_OK;
}

extern "C" int b200_rowpart_destroy(b200_rowpart *p)
{
   delete p;
   return B200_OK;
}

extern "C" int b200_rowpart_sizes(const b200_rowpart *p, int *n_rows, int *n_cols_diag, int *n_cols_offd,
                                  int *nnz_diag, int *nnz_offd, int *num_sends, int *n_send_elmts, int *num_recvs)
{
   ARG_CHECK(p != nullptr);
   if (n_rows)       { *n_rows = p->n_rows; }
   if (n_cols_diag)  { *n_cols_diag = p->n_cols_diag; }
   if (n_cols_offd)  { *n_cols_offd = p->n_cols_offd; }
   if (nnz_diag)     { *nnz_diag = (int)p->diag_j.size(); }
   if (nnz_offd)     { *nnz_offd = (int)p->offd_j.size(); }
   if (num_sends)    { *num_sends = (int)p->send_procs.size(); }
   if (n_send_elmts) { *n_send_elmts = (int)p->send_map_elmts.size(); }
   if (num_recvs)    { *num_recvs = (int)p->recv_procs.size(); }
   return B200_OK;
}

// which: 0 diag_i, 1 diag_j, 2 offd_i, 3 offd_j, 4 col_map_offd, 5 send_procs, 6 send_map_starts,
//        7 send_map_elmts, 8 recv_procs, 9 recv_vec_starts
extern "C" int b200_rowpart_int_array(const b200_rowpart *p, int which, const int **data, int *len)
{
   ARG_CHECK(p != nullptr && data != nullptr && len != nullptr);
   const std::vector<int> *v = nullptr;
   switch (which)
   {
      case 0: v = &p->diag_i; break;          case 1: v = &p->diag_j; break;
      case 2: v = &p->offd_i; break;          case 3: v = &p->offd_j; break;
      case 4: v = &p->col_map_offd; break;    case 5: v = &p->send_procs; break;
      case 6: v = &p->send_map_starts; break; case 7: v = &p->send_map_elmts; break;
      case 8: v = &p->recv_procs; break;      case 9: v = &p->recv_vec_starts; break;
      default: b200_set_error("b200_rowpart_int_array: which=%d", which); return B200_ERR_ARG;
   }
   *data = v->data(); *len = (int)v->size();
   return B200_OK;
}

// which: 0 diag values, 1 offd values
extern "C" int b200_rowpart_real_array(const b200_rowpart *p, int which, const double **data, int *len)
{
   ARG_CHECK(p != nullptr && data != nullptr && len != nullptr && (which == 0 || which == 1));
   const std::vector<double> &v = which == 0 ? p->diag_a : p->offd_a;
   *data = v.data(); *len = (int)v.size();
   return B200_OK;
}

extern "C" int b200_parcsr_create_from_rowpart(b200_ctx *ctx, const b200_rowpart *p, b200_parcsr **A)
{
   ARG_CHECK(ctx != nullptr && p != nullptr && A != nullptr);
   ARG_CHECK(p->rank == ctx->rank && p->nranks == ctx->nranks);
   const int ns = (int)p->send_procs.size(), nr = (int)p->recv_procs.size();
   return b200_parcsr_create(ctx, p->n_rows, p->n_cols_diag, p->diag_i.data(), p->diag_j.data(), p->diag_a.data(),
                             p->n_cols_offd, p->offd_i.data(), p->offd_j.data(), p->offd_a.data(),
                             ns, p->send_procs.data(), p->send_map_starts.data(), p->send_map_elmts.data(),
                             nr, p->recv_procs.data(), p->recv_vec_starts.data(), A);
}
