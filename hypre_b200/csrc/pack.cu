// pack.cu -- host-side construction of the HBM layouts of one CSR block (no CUDA calls):
// layout choice from the row-length statistics, sliced ELL / SELL-32-256 packing, and the
// lossless compressed form (16-bit column offsets + value dictionary) read by rowsum_sellc
// (spmv.cuh).  dmat_upload (matrix.cu) copies the result to the device; packed_matvec_host
// decodes it on the host with the kernels' addressing (tests of the storage format only --
// no solve path calls it).
//
// The block is the reference's hypre_CSRMatrix (src/seq_mv/csr_matrix.h: i, j, data); row
// order, entry order inside a row and the values are kept, only the encoding changes.
#include "common.cuh"
#include <algorithm>
#include <climits>
#include <cstring>
#ifdef _OPENMP
#include <omp.h>
#endif

static int g_layout_override = -1;
static int g_lanes_override  = -1;
// compressed thread-per-row storage: -1 automatic (default; B200_COMPRESS=0 in the environment turns it
// off), 0 off, CFMT_VAL8 / CFMT_VAL16 / CFMT_VAL64 force that value format where the dictionary allows it,
// CFMT_PAT8 forces the pattern storage where the block qualifies (automatic choice otherwise)
static int env_int(const char *name, int dflt) { const char *e = getenv(name); return e ? atoi(e) : dflt; }
static int g_compress = env_int("B200_COMPRESS", -1);
// VAL* storage without padding markers (rowsum_sellc_z): B200_ZPAD
static int g_zpad = env_int("B200_ZPAD", 1) != 0;
// launch options of the thread-per-row kernels (rowlaunch.cuh)
int g_rows_chunked = env_int("B200_CHUNKED", 1);
int g_pat_regs = env_int("B200_PAT_REGS", 0) != 0;
int g_rows_waves = std::min(4, std::max(1, env_int("B200_PAT_WAVES", 4)));    // grid of the pattern kernel in waves of 148 x 8 blocks

extern "C" int b200_set_compression(int mode, int no_padding_markers)
{
   ARG_CHECK(mode >= -1 && mode <= CFMT_PAT8);
   ARG_CHECK(no_padding_markers >= -1 && no_padding_markers <= 1);
   g_compress = mode;
   if (no_padding_markers >= 0) { g_zpad = no_padding_markers; }
   return B200_OK;
}

extern "C" int b200_set_kernel_options(int chunked, int pat_regs)
{
   // chunked 2: also the VAL* / plain kernels; pat_regs: bit 0 = ignored (selected register-heavier instantiations
   // that were measured slower and removed), bits 1-3 = grid of the pattern kernel in waves of resident blocks (0 leaves it)
   ARG_CHECK(chunked >= -1 && chunked <= 2 && pat_regs >= -1 && pat_regs <= 9);
   if (chunked >= 0) { g_rows_chunked = chunked; }
   if (pat_regs >= 0)
   {
      g_pat_regs = pat_regs & 1;
      if (pat_regs >> 1) { g_rows_waves = pat_regs >> 1; }
   }
   return B200_OK;
}

extern "C" int b200_set_layout_override(int layout, int lanes_per_row)
{
   ARG_CHECK(layout >= -1 && layout <= 2);
   ARG_CHECK(lanes_per_row == -1 || lanes_per_row == 1 || lanes_per_row == 2 || lanes_per_row == 4 || lanes_per_row == 8 ||
             lanes_per_row == 16 || lanes_per_row == 32);
   g_layout_override = layout;
   g_lanes_override  = lanes_per_row;
   return B200_OK;
}

// ---------------------------------------------------------------- value dictionary
// Distinct bit patterns of a[], at most `cap`; open addressing on the bit pattern (the common case --
// a handful of values -- stays in cache).  Returns false as soon as there are more than cap.
struct ValueDict {
   std::vector<double> dict;
   std::vector<unsigned long long> keys;
   std::vector<int> codes;
   static size_t slot(unsigned long long bits, size_t mask) { return (size_t)((bits * 0x9E3779B97F4A7C15ull) >> 40) & mask; }
   bool build(const std::vector<double> &a, size_t cap)
   {
      const size_t tab = 1u << 18, mask = tab - 1;     // cap <= 65536: load factor <= 1/4
      keys.assign(tab, 0); codes.assign(tab, -1);
      dict.clear();
      for (size_t k = 0; k < a.size(); k++)
      {
         unsigned long long bits; memcpy(&bits, &a[k], 8);
         size_t hsh = slot(bits, mask);
         while (codes[hsh] >= 0 && keys[hsh] != bits) { hsh = (hsh + 1) & mask; }
         if (codes[hsh] < 0)
         {
            if (dict.size() >= cap) { return false; }
            keys[hsh] = bits; codes[hsh] = (int)dict.size(); dict.push_back(a[k]);
         }
      }
      return true;
   }
   int code(double v) const
   {
      const size_t mask = keys.size() - 1;
      unsigned long long bits; memcpy(&bits, &v, 8);
      size_t hsh = slot(bits, mask);
      while (codes[hsh] >= 0 && keys[hsh] != bits) { hsh = (hsh + 1) & mask; }
      return codes[hsh];
   }
};

// ---------------------------------------------------------------- pattern dictionary (CFMT_PAT8)
// (column - row, value bits) -> code, at most PAT_CAP patterns; open addressing in a small table
constexpr int PAT_CAP = 254;            // 0xFF marks "differs between lanes" in a slice header
struct PatDict {
   static constexpr unsigned TAB = 1024;
   int           dl[TAB];
   unsigned long long vb[TAB];
   short         cd[TAB];
   std::vector<int> deltas;
   std::vector<unsigned long long> vals;
   PatDict() { for (unsigned k = 0; k < TAB; k++) { cd[k] = -1; } }
   static unsigned slot(int d, unsigned long long b)
   {
      return (unsigned)(((b ^ ((unsigned long long)(unsigned)d * 0x9E3779B97F4A7C15ull)) * 0xD6E8FEB86659FD93ull) >> 54);
   }
   int find(int d, unsigned long long b) const
   {
      unsigned h = slot(d, b);
      while (cd[h] >= 0 && (dl[h] != d || vb[h] != b)) { h = (h + 1) & (TAB - 1); }
      return cd[h];
   }
   // code of the pattern, added if new; -1 when the dictionary is full
   int add(int d, unsigned long long b)
   {
      unsigned h = slot(d, b);
      while (cd[h] >= 0 && (dl[h] != d || vb[h] != b)) { h = (h + 1) & (TAB - 1); }
      if (cd[h] >= 0) { return cd[h]; }
      if ((int)deltas.size() >= PAT_CAP) { return -1; }
      dl[h] = d; vb[h] = b; cd[h] = (short)deltas.size();
      deltas.push_back(d); vals.push_back(b);
      return cd[h];
   }
};

static inline unsigned long long dbits(double v) { unsigned long long b; memcpy(&b, &v, 8); return b; }

// Pattern storage of a thread-per-row block (format: rowsum_pat, spmv.cuh).  order: storage position -> row
// (layout 2) or nullptr; width[s]: entries of the longest row of slice s.  Returns false -- leaving P
// untouched -- when the block does not qualify: more than PAT_CAP patterns, an empty row (its padding would
// have no column of its own to gather), or a slice wider than 1020 entries.
static bool pack_patterns(const HostCsr &h, const std::vector<int> *order, int n_slices,
                          const std::vector<long long> &off_orig, PackedDMat *P)
{
   const int n = h.n_rows;
   auto row_at = [&](int p) { return order ? (*order)[(size_t)p] : p; };
   std::vector<int> w4((size_t)n_slices);
   for (int s = 0; s < n_slices; s++)
   {
      const long long w = (off_orig[(size_t)s + 1] - off_orig[(size_t)s]) / 32;
      if (w > 1020) { return false; }
      w4[(size_t)s] = (int)((w + 3) & ~3LL);
   }
   // ---- dictionary: per-thread over row ranges, merged in thread order (deterministic)
   int nthreads = 1;
#ifdef _OPENMP
   nthreads = omp_get_max_threads();
#endif
   std::vector<PatDict> local((size_t)nthreads);
   std::vector<int> slice_of_row;
   if (order)
   {
      slice_of_row.assign((size_t)n, 0);
      for (int p = 0; p < n_slices * 32; p++) { const int r = (*order)[(size_t)p]; if (r < n) { slice_of_row[(size_t)r] = p >> 5; } }
   }
   bool fail = false;
#pragma omp parallel num_threads(nthreads)
   {
      int tid = 0, nt = 1;
#ifdef _OPENMP
      tid = omp_get_thread_num(); nt = omp_get_num_threads();
#endif
      PatDict &d = local[(size_t)tid];
      const int r0 = (int)((long long)n * tid / nt), r1 = (int)((long long)n * (tid + 1) / nt);
      bool bad = false;
      for (int r = r0; r < r1 && !bad; r++)
      {
         const int beg = h.i[r], len = h.i[r + 1] - beg;
         if (len == 0) { bad = true; break; }
         for (int k = 0; k < len; k++) { if (d.add(h.j[beg + k] - r, dbits(h.a[beg + k])) < 0) { bad = true; break; } }
         const int sl = order ? slice_of_row[(size_t)r] : (r >> 5);
         if (!bad && len < w4[(size_t)sl] && d.add(h.j[beg + len - 1] - r, 0ull) < 0) { bad = true; }
      }
      if (bad)
      {
#pragma omp atomic write
         fail = true;
      }
   }
   if (fail) { return false; }
   PatDict dict;
   for (auto &d : local)
   {
      for (size_t k = 0; k < d.deltas.size(); k++) { if (dict.add(d.deltas[k], d.vals[k]) < 0) { return false; } }
   }
   // dead lanes of the last slice repeat the padding pattern of the last row (same clamped row in the kernel)
   const int last_beg = h.i[n - 1], last_len = h.i[n] - last_beg;
   const int dead_code = dict.add(h.j[last_beg + last_len - 1] - (n - 1), 0ull);
   if (dead_code < 0) { return false; }

   // ---- codes of every (slice, k, lane), entry-major per slice
   std::vector<long long> coff((size_t)n_slices + 1, 0);
   for (int s = 0; s < n_slices; s++) { coff[(size_t)s + 1] = coff[(size_t)s] + (long long)w4[(size_t)s] * 32; }
   std::vector<unsigned char> codes((size_t)coff[(size_t)n_slices]);
   std::vector<long long> bytes((size_t)n_slices + 1, 0);
#pragma omp parallel for schedule(static)
   for (int s = 0; s < n_slices; s++)
   {
      unsigned char *cs = codes.data() + coff[(size_t)s];
      const int w = w4[(size_t)s];
      for (int l = 0; l < 32; l++)
      {
         const int r = row_at(s * 32 + l);
         if (r >= n) { for (int k = 0; k < w; k++) { cs[k * 32 + l] = (unsigned char)dead_code; } continue; }
         const int beg = h.i[r], len = h.i[r + 1] - beg;
         for (int k = 0; k < len; k++) { cs[k * 32 + l] = (unsigned char)dict.find(h.j[beg + k] - r, dbits(h.a[beg + k])); }
         if (len < w)
         {
            const unsigned char pc = (unsigned char)dict.find(h.j[beg + len - 1] - r, 0ull);
            for (int k = len; k < w; k++) { cs[k * 32 + l] = pc; }
         }
      }
      int lane_words = 0;
      for (int i = 0; i < w / 4; i++)
      {
         bool mixed = false;
         for (int k = 4 * i; k < 4 * i + 4 && !mixed; k++)
         {
            for (int l = 1; l < 32; l++) { if (cs[k * 32 + l] != cs[k * 32]) { mixed = true; break; } }
         }
         lane_words += mixed;
      }
      bytes[(size_t)s + 1] = lane_words;       // 32-lane-word groups of this slice
   }
   for (int s = 0; s < n_slices; s++) { bytes[(size_t)s + 1] += bytes[(size_t)s]; }
   if (bytes[(size_t)n_slices] >= (1LL << 22)) { return false; }     // the group offset has 22 bits in word 0
   int nw_max = 0;
   for (int s = 0; s < n_slices; s++) { nw_max = std::max(nw_max, w4[(size_t)s] / 4); }
   const int stride16 = (4 + 4 * nw_max + 15) / 16;                   // 16-byte units per slice header
   P->pat_stride = stride16;
   P->pat.assign((size_t)n_slices * 16 * (size_t)stride16, 0);
   P->pat_lanes.assign((size_t)bytes[(size_t)n_slices] * 32, 0u);
#pragma omp parallel for schedule(static)
   for (int s = 0; s < n_slices; s++)
   {
      const unsigned char *cs = codes.data() + coff[(size_t)s];
      const int nw = w4[(size_t)s] / 4;
      unsigned *w0 = reinterpret_cast<unsigned *>(P->pat.data() + (size_t)s * 16 * (size_t)stride16);
      unsigned *lanew = P->pat_lanes.data() + (size_t)bytes[(size_t)s] * 32;
      const int w_true = (int)((off_orig[(size_t)s + 1] - off_orig[(size_t)s]) / 32);      // entries of the longest row
      const int last_cnt = nw > 0 ? w_true - 4 * (nw - 1) : 1;
      w0[0] = (unsigned)nw | ((unsigned)(last_cnt - 1) << 8) | ((unsigned)bytes[(size_t)s] << 10);
      for (int i = 0; i < nw; i++)
      {
         unsigned hw = 0, mixed_any = 0;
         for (int j = 0; j < 4; j++)
         {
            const unsigned char *ck = cs + (4 * i + j) * 32;
            bool mixed = false;
            for (int l = 1; l < 32; l++) { if (ck[l] != ck[0]) { mixed = true; break; } }
            hw |= (mixed ? 0xFFu : (unsigned)ck[0]) << (8 * j);
            mixed_any |= mixed;
         }
         w0[1 + i] = hw;
         if (mixed_any)
         {
            for (int l = 0; l < 32; l++)
            {
               unsigned lw = 0;
               for (int j = 0; j < 4; j++) { lw |= (unsigned)cs[(4 * i + j) * 32 + l] << (8 * j); }
               lanew[l] = lw;
            }
            lanew += 32;
         }
      }
   }
   P->dict.assign(2 * dict.deltas.size(), 0.0);
   for (size_t k = 0; k < dict.deltas.size(); k++)
   {
      memcpy(&P->dict[2 * k], &dict.vals[k], 8);
      const long long dl = (long long)(unsigned)dict.deltas[k];       // delta in the low word of the second double
      memcpy(&P->dict[2 * k + 1], &dl, 8);
   }
   P->sl_off.clear();
   P->cfmt = CFMT_PAT8;
   P->padded = coff[(size_t)n_slices];
   P->stored_bytes = (long long)P->pat.size() + 4 * (long long)P->pat_lanes.size() + 16 * (long long)dict.deltas.size();
   return true;
}

// ---------------------------------------------------------------- layout choice + packing
// Thread-per-row layouts: slice = 32 consecutive storage positions; order[p] = row stored at position p
// (>= n: padding lane).  Layout 1 keeps the row order, layout 2 sorts the rows of every 256-row window
// by decreasing length (stable) before cutting slices.
static void slice_offsets(const HostCsr &h, const std::vector<int> *order, int n_slices, std::vector<long long> *off)
{
   const int n = h.n_rows;
   std::vector<int> width((size_t)n_slices, 0);
#pragma omp parallel for schedule(static)
   for (int sidx = 0; sidx < n_slices; sidx++)
   {
      int w = 0;
      for (int l = 0; l < 32; l++)
      {
         const int p = sidx * 32 + l;
         const int r = order ? (*order)[(size_t)p] : p;
         if (r < n) { w = std::max(w, h.i[r + 1] - h.i[r]); }
      }
      width[(size_t)sidx] = w;
   }
   off->assign((size_t)n_slices + 1, 0);
   for (int sidx = 0; sidx < n_slices; sidx++) { (*off)[(size_t)sidx + 1] = (*off)[(size_t)sidx] + (long long)width[(size_t)sidx] * 32; }
}

static void sorted_order(const HostCsr &h, int n_slices, std::vector<int> *sorted)
{
   const int n = h.n_rows, np = n_slices * 32;
   sorted->resize((size_t)np);
#pragma omp parallel for schedule(static)
   for (int w0 = 0; w0 < np; w0 += 256)
   {
      const int w1 = std::min(w0 + 256, np);
      for (int p = w0; p < w1; p++) { (*sorted)[(size_t)p] = p; }
      std::stable_sort(sorted->begin() + w0, sorted->begin() + w1, [&](int a, int b) {
         const int la = a < n ? h.i[a + 1] - h.i[a] : -1, lb = b < n ? h.i[b + 1] - h.i[b] : -1;
         return la > lb;
      });
   }
}

void dmat_pack(const HostCsr &h, int allow_sell, PackedDMat *P)
{
   *P = PackedDMat();
   const int n = h.n_rows;
   const long long nnz = n > 0 ? (long long)h.i[n] : 0;
   if (n == 0) { return; }

   int max_row = 0;
   for (int r = 0; r < n; r++) { max_row = std::max(max_row, h.i[r + 1] - h.i[r]); }
   P->max_row = max_row;
   P->avg_row = (double)nnz / (double)n;
   int lanes;
   if      (P->avg_row > 48.0) { lanes = 32; }
   else if (P->avg_row > 24.0) { lanes = 16; }
   else if (P->avg_row > 12.0) { lanes = 8; }
   else if (P->avg_row >  5.0) { lanes = 4; }
   else if (P->avg_row >  3.0) { lanes = 2; }
   else                        { lanes = 1; }   // interpolation-like rows (1-4 entries): thread per row
   if (g_lanes_override > 0) { lanes = g_lanes_override; }
   P->lanes = lanes;

   // thread-per-row layouts need rows to fill the machine (148 SMs x 2048 threads) and padding of at
   // most 12.5 %; smaller blocks (coarse levels) keep CSR with several lanes per row
   const int n_slices = (n + 31) / 32;
   const double nnz_d = (double)nnz;
   std::vector<int> sorted;
   std::vector<long long> off_orig;
   int layout = 0;
   const bool forced = g_layout_override >= 0 && allow_sell;
   if (forced) { layout = g_layout_override; }
   else if (allow_sell && nnz > 0 && n >= 131072)
   {
      slice_offsets(h, nullptr, n_slices, &off_orig);
      if ((double)off_orig[(size_t)n_slices] <= 1.125 * nnz_d) { layout = 1; }
      else
      {
         sorted_order(h, n_slices, &sorted);
         slice_offsets(h, &sorted, n_slices, &off_orig);
         if ((double)off_orig[(size_t)n_slices] <= 1.125 * nnz_d) { layout = 2; }
      }
   }
   P->layout = layout;
   if (layout == 0) { return; }
   if (forced)
   {
      if (layout == 2) { sorted_order(h, n_slices, &sorted); }
      slice_offsets(h, layout == 2 ? &sorted : nullptr, n_slices, &off_orig);
   }
   const std::vector<int> *order = layout == 2 ? &sorted : nullptr;
   auto row_at = [&](int p) { return order ? (*order)[(size_t)p] : p; };

   // ---- compressed storage (rowsum_sellc): pick the value format, round the slice widths up to whole
   // chunks of 4 entries, and keep it only if a pass over the block then reads fewer bytes
   ValueDict vd;
   int cfmt = CFMT_NONE;
   bool zpad = false;
   std::vector<long long> sl_off = off_orig;
   // pattern storage first: where a block qualifies (stencils and their first Galerkin products) nothing else
   // comes close -- one byte per entry at most, none for lane-uniform entries
   if (g_compress == -1 || g_compress == CFMT_PAT8)
   {
      PackedDMat Q = *P;
      if (pack_patterns(h, order, n_slices, off_orig, &Q))
      {
         Q.n_slices = n_slices;
         if (layout == 2)
         {
            Q.perm.resize(sorted.size());
            Q.h_pos.assign((size_t)n, 0);
            for (size_t p = 0; p < sorted.size(); p++)
            {
               Q.perm[p] = (unsigned short)(sorted[p] - (int)(p / 256) * 256);
               if (sorted[p] < n) { Q.h_pos[(size_t)sorted[p]] = (int)p; }
            }
         }
         *P = std::move(Q);
         return;
      }
   }
   if (g_compress != 0)
   {
      cfmt = CFMT_VAL64;
      // storage without padding markers (rowsum_sellc_z): padding lanes gather a valid column and
      // multiply it by a dictionary entry 0.0 (as the plain sliced-ELL layout does); one code is kept for it
      zpad = g_zpad != 0;
      const int vmode = g_compress == CFMT_PAT8 ? -1 : g_compress;
      if (vmode != CFMT_VAL64 && vd.build(h.a, (vmode == CFMT_VAL8 ? 256 : 65536) - (size_t)zpad))
      {
         cfmt = (vd.dict.size() + (size_t)zpad <= 256 && vmode != CFMT_VAL16) ? CFMT_VAL8 : CFMT_VAL16;
      }
      std::vector<long long> off4((size_t)n_slices + 1, 0);
      for (int sidx = 0; sidx < n_slices; sidx++)
      {
         const long long w = (off_orig[(size_t)sidx + 1] - off_orig[(size_t)sidx]) / 32;
         off4[(size_t)sidx + 1] = off4[(size_t)sidx] + ((w + 3) & ~3LL) * 32;
      }
      const long long per_entry = 2 + (cfmt == CFMT_VAL8 ? 1 : cfmt == CFMT_VAL16 ? 2 : 8);
      if (vmode > 0 || per_entry * off4[(size_t)n_slices] + off4[(size_t)n_slices] / 8 < 12 * off_orig[(size_t)n_slices])
      {
         sl_off.swap(off4);
      }
      else { cfmt = CFMT_NONE; }
   }
   const long long padded = sl_off[(size_t)n_slices];
   P->n_slices = n_slices; P->padded = padded; P->cfmt = cfmt;

   if (layout == 2)
   {
      P->perm.resize(sorted.size());
      P->h_pos.assign((size_t)n, 0);
      for (size_t p = 0; p < sorted.size(); p++)
      {
         P->perm[p] = (unsigned short)(sorted[p] - (int)(p / 256) * 256);    // < 256 + 32
         if (sorted[p] < n) { P->h_pos[(size_t)sorted[p]] = (int)p; }
      }
   }

   // entry-major arrays ([slice][k][lane]): the plain layout, the fp64 values of CFMT_VAL64 and the
   // 32-bit columns kept for escape groups.  Padding gathers a column the row already touches, value 0.
   auto fill_entry_major = [&](bool want_ci, bool want_va) {
      if (!want_ci && !want_va) { return; }
      if (want_ci) { P->sci.resize((size_t)padded); }
      if (want_va) { P->sva.resize((size_t)padded); }
#pragma omp parallel for schedule(static)
      for (int sidx = 0; sidx < n_slices; sidx++)
      {
         const long long off = sl_off[(size_t)sidx];
         const int w = (int)((sl_off[(size_t)sidx + 1] - off) / 32);
         for (int l = 0; l < 32; l++)
         {
            const int r = row_at(sidx * 32 + l);
            const int beg = r < n ? h.i[r] : 0, len = r < n ? h.i[r + 1] - h.i[r] : 0;
            const int padc = len > 0 ? h.j[beg + len - 1] : 0;
            for (int k = 0; k < w; k++)
            {
               const size_t p = (size_t)off + (size_t)k * 32 + (size_t)l;
               if (want_ci) { P->sci[p] = k < len ? h.j[beg + k] : padc; }
               if (want_va) { P->sva[p] = k < len ? h.a[beg + k] : 0.0; }
            }
         }
      }
   };
   if (cfmt == CFMT_NONE)
   {
      fill_entry_major(true, true);
      P->stored_bytes = 12 * padded + 8 * (long long)n_slices;
      P->sl_off.swap(sl_off);
      return;
   }

   // chunk-transposed arrays ([slice][k/4][lane][k%4]): 16-bit column offsets from the base of
   // (slice, k), value codes; CFMT_PAD marks padding lanes
   const size_t n_groups = (size_t)(padded / 32);
   const size_t csz = cfmt == CFMT_VAL8 ? 1 : cfmt == CFMT_VAL16 ? 2 : 0;
   P->base.assign(n_groups, 0);
   int zero_code = 0;
   if (csz && zpad)
   {
      zero_code = vd.dict.empty() ? -1 : vd.code(0.0);
      if (zero_code < 0) { zero_code = (int)vd.dict.size(); vd.dict.push_back(0.0); }    // codes of h.a are assigned already
   }
   P->zpad = zpad;
   P->idx.assign((size_t)padded, zpad ? (unsigned short)0 : CFMT_PAD);
   if (csz == 1) { P->code.assign((size_t)padded, (unsigned char)zero_code); }
   else
   {
      P->code.assign((size_t)padded * csz, 0);
      if (csz == 2 && zero_code != 0)
      {
         const unsigned short z16 = (unsigned short)zero_code;
#pragma omp parallel for schedule(static)
         for (long long p = 0; p < padded; p++) { memcpy(&P->code[(size_t)p * 2], &z16, 2); }
      }
   }
   if (csz && vd.dict.empty()) { vd.dict.push_back(0.0); }
   long long n_esc = 0;
#pragma omp parallel for schedule(static) reduction(+ : n_esc)
   for (int sidx = 0; sidx < n_slices; sidx++)
   {
      const long long off = sl_off[(size_t)sidx];
      const int w = (int)((sl_off[(size_t)sidx + 1] - off) / 32);
      int beg[32], len[32];
      for (int l = 0; l < 32; l++)
      {
         const int r = row_at(sidx * 32 + l);
         beg[l] = r < n ? h.i[r] : 0; len[l] = r < n ? h.i[r + 1] - h.i[r] : 0;
      }
      for (int k = 0; k < w; k++)
      {
         int lo = INT_MAX, hi = INT_MIN;
         for (int l = 0; l < 32; l++) { if (k < len[l]) { const int c = h.j[beg[l] + k]; lo = std::min(lo, c); hi = std::max(hi, c); } }
         int padcol[32];
         if (zpad)
         {
            // a padding lane multiplies 0.0 with x at one of ITS OWN row's columns (so that nothing a row does
            // not reference can reach it): the one that widens [lo, hi] least.  Rows without entries take lo.
            for (int l = 0; l < 32; l++)
            {
               padcol[l] = -1;
               if (k < len[l] || len[l] == 0) { continue; }
               long long best = LLONG_MAX;
               for (int q = 0; q < len[l]; q++)
               {
                  const int c = h.j[beg[l] + q];
                  const long long span = lo > hi ? 0 : (long long)std::max(hi, c) - (long long)std::min(lo, c);
                  if (span < best) { best = span; padcol[l] = c; }
               }
               lo = std::min(lo, padcol[l]); hi = std::max(hi, padcol[l]);
            }
         }
         if (lo > hi) { continue; }                    // nothing live: base 0, all lanes CFMT_PAD (or offset 0 x 0.0)
         const bool esc = (long long)hi - (long long)lo > 65534;
         P->base[(size_t)(off / 32) + (size_t)k] = esc ? CFMT_ESC : lo;
         n_esc += esc;
         for (int l = 0; l < 32; l++)
         {
            if (k >= len[l])
            {
               if (zpad && !esc && padcol[l] >= 0)
               {
                  P->idx[(size_t)off + (size_t)(k >> 2) * 128 + (size_t)l * 4 + (size_t)(k & 3)] = (unsigned short)(padcol[l] - lo);
               }
               continue;
            }
            const size_t p = (size_t)off + (size_t)(k >> 2) * 128 + (size_t)l * 4 + (size_t)(k & 3);
            P->idx[p] = esc ? (unsigned short)0 : (unsigned short)(h.j[beg[l] + k] - lo);
            if (csz)
            {
               const int cd = vd.code(h.a[beg[l] + k]);
               if (csz == 1) { P->code[p] = (unsigned char)cd; }
               else { const unsigned short c16 = (unsigned short)cd; memcpy(&P->code[p * 2], &c16, 2); }
            }
         }
      }
   }
   P->n_esc = n_esc;
   fill_entry_major(n_esc > 0, cfmt == CFMT_VAL64);
   if (csz) { P->dict.swap(vd.dict); }
   const long long per_entry = 2 + (cfmt == CFMT_VAL8 ? 1 : cfmt == CFMT_VAL16 ? 2 : 8);
   P->stored_bytes = per_entry * padded + 4 * (long long)n_groups + 8 * (long long)n_slices + 4 * 32 * n_esc;
   P->sl_off.swap(sl_off);
}

// ---------------------------------------------------------------- host decode (tests of the format)
// y = A x from the packed arrays, with the addressing of rows_range / rowsum_sell / rowsum_sellc
// (spmv.cuh) and the same fma chain per row.  Layout 0 has no packed form (the CSR arrays are used).
int packed_matvec_host(const PackedDMat &P, const HostCsr &h, const double *x, double *y)
{
   const int n = h.n_rows;
   if (P.layout == 0)
   {
      for (int r = 0; r < n; r++)
      {
         double t = 0.0;
         for (int k = h.i[r]; k < h.i[r + 1]; k++) { t = fma(h.a[k], x[h.j[k]], t); }
         y[r] = t;
      }
      return B200_OK;
   }
   for (int slice = 0; slice < P.n_slices; slice++)
   {
      const long long off = P.cfmt == CFMT_PAT8 ? 0 : P.sl_off[(size_t)slice];
      const long long span = P.cfmt == CFMT_PAT8 ? 0 : P.sl_off[(size_t)slice + 1] - off;
      for (int lane = 0; lane < 32; lane++)
      {
         int row = slice * 32 + lane;
         if (!P.perm.empty()) { row = (slice >> 3) * 256 + P.perm[(size_t)row]; }
         double t = 0.0;
         bool stray = false;      // a row beyond n_rows multiplied a non-zero value
         if (P.cfmt == CFMT_PAT8)
         {
            // the addressing of rowsum_pat: fixed-stride header, lane words in their own stream
            const unsigned *w0 = reinterpret_cast<const unsigned *>(P.pat.data() + (size_t)slice * 16 * (size_t)P.pat_stride);
            const unsigned nw = w0[0] & 0xFFu;
            if (4 + 4 * (int)nw > 16 * P.pat_stride) { b200_set_error("pattern header wider than its stride"); return B200_ERR_ARG; }
            const unsigned *lp = P.pat_lanes.data() + (size_t)(w0[0] >> 10) * 32 + lane;
            const int last_cnt = (int)((w0[0] >> 8) & 3u) + 1;
            const int rowc = row < n ? row : n - 1;
            for (unsigned i = 0; i < nw; i++)
            {
               unsigned hw = w0[1 + i], m = 0;
               for (int j = 0; j < 4; j++) { if (((hw >> (8 * j)) & 0xFFu) == 0xFFu) { m |= 0xFFu << (8 * j); } }
               if (m)
               {
                  if (lp >= P.pat_lanes.data() + P.pat_lanes.size()) { b200_set_error("pattern stream overrun"); return B200_ERR_ARG; }
                  hw = (*lp & m) | (hw & ~m); lp += 32;
               }
               for (int j = 0; j < (i + 1 == nw ? last_cnt : 4); j++)
               {
                  const unsigned cd = (hw >> (8 * j)) & 0xFFu;
                  if (2 * (size_t)cd + 1 >= P.dict.size()) { b200_set_error("pattern code out of range"); return B200_ERR_ARG; }
                  const double v = P.dict[2 * cd];
                  int delta; memcpy(&delta, &P.dict[2 * cd + 1], 4);
                  const int col = rowc + delta;
                  if (col < 0 || col >= h.n_cols) { b200_set_error("packed column out of range"); return B200_ERR_ARG; }
                  t = fma(v, x[col], t);
                  stray |= (row >= n && v != 0.0);
               }
            }
         }
         else if (P.cfmt == CFMT_NONE)
         {
            const int width = (int)(span >> 5);
            for (int k = 0; k < width; k++)
            {
               const size_t p = (size_t)off + (size_t)k * 32 + (size_t)lane;
               const int c = P.sci[p];
               if (c < 0 || c >= h.n_cols) { b200_set_error("packed column out of range"); return B200_ERR_ARG; }
               t = fma(P.sva[p], x[c], t);
               stray |= (row >= n && P.sva[p] != 0.0);
            }
         }
         else
         {
            if (span & 127) { b200_set_error("compressed slice width not a multiple of 4"); return B200_ERR_ARG; }
            const int n_chunks = (int)(span >> 7);
            for (int c = 0; c < n_chunks; c++)
            {
               for (int j = 0; j < 4; j++)
               {
                  const size_t p = (size_t)off + (size_t)c * 128 + (size_t)lane * 4 + (size_t)j;
                  const unsigned o = P.idx[p];
                  if (o == CFMT_PAD) { continue; }
                  const int b = P.base[(size_t)(off >> 5) + (size_t)c * 4 + (size_t)j];
                  const size_t e = (size_t)off + (size_t)(c * 4 + j) * 32 + (size_t)lane;      // entry-major position
                  int col;
                  if (b == CFMT_ESC)
                  {
                     if (P.sci.empty()) { b200_set_error("escape group without 32-bit columns"); return B200_ERR_ARG; }
                     col = P.sci[e];
                  }
                  else { col = b + (int)o; }
                  if (col < 0 || col >= h.n_cols) { b200_set_error("packed column out of range"); return B200_ERR_ARG; }
                  double v;
                  if (P.cfmt == CFMT_VAL8)       { v = P.dict[P.code[p]]; }
                  else if (P.cfmt == CFMT_VAL16) { unsigned short c16; memcpy(&c16, &P.code[p * 2], 2); v = P.dict[c16]; }
                  else                           { v = P.sva[e]; }
                  t = fma(v, x[col], t);
                  stray |= (row >= n && v != 0.0);
               }
            }
         }
         if (row < n) { y[row] = t; }
         else if (stray) { b200_set_error("padding lane carries entries"); return B200_ERR_ARG; }
      }
   }
   return B200_OK;
}

// C-ABI: pack a CSR block with the current layout / compression settings and decode it on the host.
extern "C" int b200_storage_decode_host(int n_rows, int n_cols, const int *ai, const int *aj, const double *aa,
                                        const double *x, double *y, int *layout, int *value_format,
                                        long long *stored_bytes, long long *n_escape_groups)
{
   ARG_CHECK(n_rows >= 0 && n_cols >= 0 && (n_rows == 0 || (ai != nullptr && x != nullptr && y != nullptr)));
   HostCsr h;
   h.n_rows = n_rows; h.n_cols = n_cols;
   h.i.assign(1, 0);
   if (n_rows > 0)
   {
      h.i.assign(ai, ai + n_rows + 1);
      h.j.assign(aj, aj + ai[n_rows]);
      h.a.assign(aa, aa + ai[n_rows]);
   }
   PackedDMat P;
   dmat_pack(h, 1, &P);
   if (layout) { *layout = P.layout; }
   if (value_format) { *value_format = P.cfmt; }
   if (stored_bytes) { *stored_bytes = P.layout >= 1 ? P.stored_bytes : 12LL * (long long)h.j.size() + 4LL * (n_rows + 1); }
   if (n_escape_groups) { *n_escape_groups = P.n_esc; }
   return packed_matvec_host(P, h, x, y);
}
