// Host side of the SpMM engine: plan construction (CSR upload, row classes, transpose), kernel
// dispatch by feature width / dtype / alignment, the host-buffer end-to-end entry point.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <thread>
#include <vector>

#include "common.h"
#include "spmm_kernels.cuh"

namespace gcnb {

struct Matrix {  // one CSR resident in HBM + its block-class work lists
  int64_t m = 0, n = 0, nnz = 0;
  int32_t* rowptr = nullptr;      // [m+1]
  int2* cv = nullptr;             // [nnz+2] interleaved (col, val-bits)
  WorkItem* items = nullptr;      // [n_items]
  int32_t* split_rows = nullptr;  // [n_split] rows cut into >1 item
  int32_t* split_ptr = nullptr;   // [n_split+1] first partial slot of each split row
  // stream class (wide feature rows): segments of ~kSegNnz non-zeros — runs of whole short rows, or
  // pieces of one long row (their partial sums meet in spmm_reduce_kernel through s_split_*)
  Seg* segs = nullptr;            // [n_segs]
  int32_t* s_split_rows = nullptr;
  int32_t* s_split_ptr = nullptr;
  int64_t n_segs = 0, s_n_split = 0, s_n_partials = 0;
  // panel class (high mean degree): items of 32-row panels + per-row block boundary table, see
  // spmm_kernels.cuh spmm_panel_kernel
  PanelItem* p_items = nullptr;
  int32_t* p_subptr = nullptr;
  int32_t* p_split_rows = nullptr;
  int32_t* p_split_ptr = nullptr;
  int64_t p_n_items = 0, p_n_split = 0, p_n_partials = 0, p_distinct = 0;
  int64_t n_items = 0, n_split = 0, n_partials = 0, n_long_rows = 0, max_row_nnz = 0;
  size_t bytes = 0;

  void release() {
    cudaFree(rowptr); cudaFree(cv); cudaFree(items); cudaFree(split_rows); cudaFree(split_ptr);
    cudaFree(segs); cudaFree(s_split_rows); cudaFree(s_split_ptr);
    cudaFree(p_items); cudaFree(p_subptr); cudaFree(p_split_rows); cudaFree(p_split_ptr);
    p_items = nullptr; p_subptr = nullptr; p_split_rows = nullptr; p_split_ptr = nullptr;
    s_split_rows = nullptr; s_split_ptr = nullptr;
    rowptr = nullptr; cv = nullptr; items = nullptr; split_rows = nullptr; split_ptr = nullptr;
    segs = nullptr;
  }
};

}  // namespace gcnb

struct gcnb_plan {
  int device = 0;
  int n_sm = 0;
  bool symmetric = false;
  bool has_t = false;
  gcnb::Matrix a, at;
  // staging for gcnb_spmm_host (grown on demand; that entry point is not re-entrant per plan)
  float* stage_b = nullptr;
  float* stage_c = nullptr;
  size_t stage_b_bytes = 0, stage_c_bytes = 0;
};

namespace gcnb {

// non-zeros per stream segment: kSegNnz on large matrices, smaller when the matrix would not give
// every SM a few dozen segments
static int64_t seg_target(int64_t nnz, int n_sm) {
  static const int64_t forced = getenv("GCNB_SEG_NNZ") ? atoll(getenv("GCNB_SEG_NNZ")) : 0;
  if (forced > 0) return forced;
  const int64_t want = nnz / ((int64_t)n_sm * 64);
  return std::max<int64_t>(64, std::min<int64_t>(kSegNnz, want));
}


// ---- panel class, plan time ------------------------------------------------------------------------
// For every panel of kPanelRows consecutive rows: cut its columns (with multiplicity) into blocks of
// ~tile_nnz non-zeros, record for every row where its non-zeros cross the cuts, cut heavy panels into
// items of ~item_nnz non-zeros along the block list.  Host threads over panels.
struct PanelBuild {
  std::vector<PanelItem> items;
  std::vector<int32_t> subptr, split_rows, split_ptr;
  int64_t n_partials = 0, distinct = 0;
};

static void build_panels(const int32_t* rowptr, const int32_t* col, int64_t m, int64_t nnz, int tile_nnz,
                         int item_nnz, PanelBuild* out) {
  const int64_t n_panels = (m + kPanelRows - 1) / kPanelRows;
  struct One { std::vector<int32_t> table; int32_t nb = 1; int64_t distinct = 0; };
  std::vector<One> per(n_panels);
  std::atomic<int64_t> next{0};
  auto work = [&]() {
    std::vector<int32_t> cols, cuts;
    for (;;) {
      const int64_t p = next.fetch_add(1);
      if (p >= n_panels) break;
      const int64_t r0 = p * kPanelRows, r1 = std::min<int64_t>(m, r0 + kPanelRows);
      const int64_t zb = rowptr[r0], ze = rowptr[r1], P = ze - zb;
      cols.assign(col + zb, col + ze);
      std::sort(cols.begin(), cols.end());
      int64_t distinct = P > 0 ? 1 : 0;
      for (int64_t i = 1; i < P; ++i) distinct += cols[i] != cols[i - 1];
      int64_t nb = std::max<int64_t>(1, std::min<int64_t>(32767, (P + tile_nnz - 1) / tile_nnz));
      cuts.clear();
      for (int64_t j = 1; j < nb; ++j) {
        const int32_t c = cols[(size_t)(j * P / nb)];
        if (cuts.empty() || c > cuts.back()) cuts.push_back(c);   // strictly increasing column cuts
      }
      nb = (int64_t)cuts.size() + 1;
      One& o = per[p];
      o.nb = (int32_t)nb;
      o.distinct = distinct;
      o.table.resize((size_t)kPanelRows * (nb + 1));
      for (int r = 0; r < kPanelRows; ++r) {
        int32_t* t = o.table.data() + (size_t)r * (nb + 1);
        if (r0 + r >= m) {
          for (int64_t j = 0; j <= nb; ++j) t[j] = (int32_t)ze;   // padding row: empty everywhere
          continue;
        }
        const int32_t* cb = col + rowptr[r0 + r];
        const int32_t* ce = col + rowptr[r0 + r + 1];
        t[0] = rowptr[r0 + r];
        const int32_t* pos = cb;
        for (int64_t j = 1; j < nb; ++j) {
          pos = std::lower_bound(pos, ce, cuts[(size_t)j - 1]);     // block j starts at the first column >= cut
          t[j] = (int32_t)(rowptr[r0 + r] + (pos - cb));
        }
        t[nb] = rowptr[r0 + r + 1];
      }
    }
  };
  unsigned nt = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
  if (getenv("GCNB_HOST_THREADS")) nt = (unsigned)std::max(1, atoi(getenv("GCNB_HOST_THREADS")));
  if (nnz < (1 << 20)) nt = 1;
  std::vector<std::thread> th;
  for (unsigned i = 1; i < nt; ++i) th.emplace_back(work);
  work();
  for (auto& t : th) t.join();

  out->split_ptr.clear();
  for (int64_t p = 0; p < n_panels; ++p) {
    const One& o = per[p];
    const int64_t r0 = p * kPanelRows, r1 = std::min<int64_t>(m, r0 + kPanelRows);
    const int64_t P = (int64_t)rowptr[r1] - rowptr[r0];
    const int32_t sub_off = (int32_t)out->subptr.size();
    out->subptr.insert(out->subptr.end(), o.table.begin(), o.table.end());
    out->distinct += o.distinct;
    const int64_t n_items = std::max<int64_t>(1, std::min<int64_t>(o.nb, (P + item_nnz - 1) / item_nnz));
    if (n_items == 1) {
      out->items.push_back({(int32_t)r0, 0, o.nb, -1, 1, sub_off, o.nb, 0});
    } else {
      const int64_t base = out->n_partials;
      for (int64_t i = 0; i < n_items; ++i)
        out->items.push_back({(int32_t)r0, (int32_t)(i * o.nb / n_items), (int32_t)((i + 1) * o.nb / n_items),
                              (int32_t)(base + i), (int32_t)n_items, sub_off, o.nb, 0});
      for (int64_t r = r0; r < r1; ++r) {
        out->split_rows.push_back((int32_t)r);
        out->split_ptr.push_back((int32_t)(base + (r - r0) * n_items));
      }
      out->n_partials += (r1 - r0) * n_items;   // padding rows of the last panel write nothing
    }
  }
  out->split_ptr.push_back((int32_t)out->n_partials);
}

static int upload_matrix(const int32_t* rowptr, const int32_t* col, const float* val, int64_t m,
                         int64_t n, int64_t nnz, int n_sm, int flags, Matrix* M) {
  M->m = m; M->n = n; M->nnz = nnz;
  // block-class work lists (host, O(m))
  std::vector<WorkItem> items;
  std::vector<int32_t> split_rows, split_ptr;
  int64_t n_partials = 0, max_len = 0, n_long = 0;
  for (int64_t r = 0; r < m; ++r) {
    const int64_t beg = rowptr[r], end = rowptr[r + 1];
    const int64_t len = end - beg;
    max_len = std::max(max_len, len);
    if (len < kLongRow) continue;
    ++n_long;
    if (len <= kChunk) {
      items.push_back({(int32_t)r, (int32_t)beg, (int32_t)end, -1});
    } else {
      // equal-sized pieces, each <= kChunk
      const int64_t pieces = (len + kChunk - 1) / kChunk;
      const int64_t per = (len + pieces - 1) / pieces;
      split_rows.push_back((int32_t)r);
      split_ptr.push_back((int32_t)n_partials);
      for (int64_t b = beg; b < end; b += per)
        items.push_back({(int32_t)r, (int32_t)b, (int32_t)std::min(end, b + per),
                         (int32_t)n_partials++});
    }
  }
  split_ptr.push_back((int32_t)n_partials);
  // stream-class segments: maximal runs of consecutive short rows, cut every ~seg_nnz non-zeros
  // (and every 4096 rows, so that runs of empty rows are spread as well); a long row closes the run
  // and is emitted as one segment or, beyond kSegLong non-zeros, as equal pieces with partial slots
  std::vector<Seg> segs;
  std::vector<int32_t> s_split_rows, s_split_ptr;
  int64_t s_n_partials = 0;
  {
    const int64_t seg_nnz = seg_target(nnz, n_sm);
    // On a small matrix (seg_nnz below kSegNnz: the remote-column block of a row partition, 1.4 M non-zeros
    // at 8 GPUs) a 1024-non-zero piece of a hub row is a serial chain of 256 ring turns, each an HBM latency:
    // it alone took 0.25 ms while the other warps were done after 0.08 (profiles/r02_af_*).  Rows are
    // therefore "long" from 2 x seg_nnz non-zeros on and cut into pieces of at most that.
    static const int short_pieces = getenv("GCNB_SHORT_PIECES") ? atoi(getenv("GCNB_SHORT_PIECES")) : 1;
    const int64_t long_row = short_pieces ? std::min<int64_t>(kLongRow, std::max<int64_t>(128, 2 * seg_nnz)) : kLongRow;
    const int64_t seg_long = short_pieces ? std::min<int64_t>(kSegLong, std::max<int64_t>(128, 2 * seg_nnz)) : kSegLong;
    int64_t r0 = -1;
    auto close_run = [&](int64_t r_end) {
      if (r0 >= 0) segs.push_back({(int32_t)r0, (int32_t)r_end, rowptr[r0], rowptr[r_end]});
      r0 = -1;
    };
    for (int64_t r = 0; r < m; ++r) {
      const int64_t beg = rowptr[r], end = rowptr[r + 1], len = end - beg;
      if (len >= long_row) {
        close_run(r);
        if (len <= seg_long) {
          segs.push_back({(int32_t)r, (int32_t)(r + 1), (int32_t)beg, (int32_t)end});
        } else {
          const int64_t pieces = (len + seg_long - 1) / seg_long;
          const int64_t per = (len + pieces - 1) / pieces;
          s_split_rows.push_back((int32_t)r);
          s_split_ptr.push_back((int32_t)s_n_partials);
          for (int64_t b = beg; b < end; b += per)
            segs.push_back({(int32_t)r, (int32_t)(-1 - s_n_partials++), (int32_t)b,
                            (int32_t)std::min(end, b + per)});
        }
        continue;
      }
      if (r0 < 0) r0 = r;
      if (end - rowptr[r0] >= seg_nnz || r + 1 - r0 >= 4096) close_run(r + 1);
    }
    close_run(m);
    s_split_ptr.push_back((int32_t)s_n_partials);
  }
  M->s_n_split = (int64_t)s_split_rows.size();
  M->s_n_partials = s_n_partials;
  M->n_segs = (int64_t)segs.size();
  M->n_items = (int64_t)items.size();
  M->n_split = (int64_t)split_rows.size();
  M->n_partials = n_partials;
  M->n_long_rows = n_long;
  M->max_row_nnz = max_len;

  GCNB_CHECK_CUDA(cudaMalloc(&M->rowptr, (size_t)(m + 1) * sizeof(int32_t)));
  GCNB_CHECK_CUDA(cudaMalloc(&M->cv, (size_t)(nnz + 2) * sizeof(int2)));
  M->bytes = (size_t)(m + 1) * 4 + (size_t)(nnz + 2) * 8;
  GCNB_CHECK_CUDA(cudaMemcpy(M->rowptr, rowptr, (size_t)(m + 1) * sizeof(int32_t),
                             cudaMemcpyHostToDevice));
  GCNB_CHECK_CUDA(cudaMemset(M->cv + nnz, 0, 2 * sizeof(int2)));
  if (nnz > 0) {
    int32_t* d_col = nullptr;
    float* d_val = nullptr;
    GCNB_CHECK_CUDA(cudaMalloc(&d_col, (size_t)nnz * 4));
    GCNB_CHECK_CUDA(cudaMalloc(&d_val, (size_t)nnz * 4));
    GCNB_CHECK_CUDA(cudaMemcpy(d_col, col, (size_t)nnz * 4, cudaMemcpyHostToDevice));
    GCNB_CHECK_CUDA(cudaMemcpy(d_val, val, (size_t)nnz * 4, cudaMemcpyHostToDevice));
    const int64_t blocks = (nnz + 255) / 256;
    interleave_kernel<<<(unsigned)blocks, 256>>>(d_col, d_val, M->cv, nnz);
    GCNB_CHECK_CUDA(cudaGetLastError());
    GCNB_CHECK_CUDA(cudaDeviceSynchronize());
    cudaFree(d_col);
    cudaFree(d_val);
  }
  if (!segs.empty()) {
    GCNB_CHECK_CUDA(cudaMalloc(&M->segs, segs.size() * sizeof(Seg)));
    GCNB_CHECK_CUDA(cudaMemcpy(M->segs, segs.data(), segs.size() * sizeof(Seg), cudaMemcpyHostToDevice));
    M->bytes += segs.size() * sizeof(Seg);
  }
  if (!s_split_rows.empty()) {
    GCNB_CHECK_CUDA(cudaMalloc(&M->s_split_rows, s_split_rows.size() * 4));
    GCNB_CHECK_CUDA(cudaMalloc(&M->s_split_ptr, s_split_ptr.size() * 4));
    GCNB_CHECK_CUDA(cudaMemcpy(M->s_split_rows, s_split_rows.data(), s_split_rows.size() * 4,
                               cudaMemcpyHostToDevice));
    GCNB_CHECK_CUDA(cudaMemcpy(M->s_split_ptr, s_split_ptr.data(), s_split_ptr.size() * 4,
                               cudaMemcpyHostToDevice));
    M->bytes += (s_split_rows.size() + s_split_ptr.size()) * 4;
  }
  // panel class: OFF unless asked for (GCNB_PLAN_PANEL / GCNB_PANEL=1).  Measured on the Reddit-shaped graph
  // (profiles/r02_b_*, r02_d_ncu_reddit_deg_*): it removes a third of the L2->SM bytes (L1 hit rate 16 % -> 42 %)
  // and is no faster than the stream class (3.17 vs 3.09 ms in degree order) because an L1 hit occupies the LSU
  // data pipe exactly like the miss it replaces; in an order that puts rows of very different length into one
  // panel it is 10x slower (35.5 vs 3.43 ms).  GCNB_PANEL: 0 never (default), 1 always, -1 = the size /
  // sharing rule below.
  {
    static const int panel_env = getenv("GCNB_PANEL") ? atoi(getenv("GCNB_PANEL")) : 0;
    const int panel_mode = (flags & GCNB_PLAN_NO_PANEL) ? 0 : (flags & GCNB_PLAN_PANEL) ? 1 : panel_env;
    static const int tile_nnz = getenv("GCNB_PANEL_TILE") ? atoi(getenv("GCNB_PANEL_TILE")) : 512;
    static const int item_nnz = getenv("GCNB_PANEL_ITEM") ? atoi(getenv("GCNB_PANEL_ITEM")) : 32768;
    static const double panel_min_share = getenv("GCNB_PANEL_MIN_SHARE") ? atof(getenv("GCNB_PANEL_MIN_SHARE")) : 1.25;
    if (panel_mode != 0 && m > 0 && nnz > 0 && (panel_mode == 1 || nnz >= 96 * m)) {
      PanelBuild pb;
      build_panels(rowptr, col, m, nnz, std::max(32, tile_nnz), std::max(1024, item_nnz), &pb);
      M->p_distinct = pb.distinct;
      if (panel_mode == 1 || (double)nnz >= panel_min_share * (double)pb.distinct) {
        M->p_n_items = (int64_t)pb.items.size();
        M->p_n_split = (int64_t)pb.split_rows.size();
        M->p_n_partials = pb.n_partials;
        GCNB_CHECK_CUDA(cudaMalloc(&M->p_items, pb.items.size() * sizeof(PanelItem)));
        GCNB_CHECK_CUDA(cudaMemcpy(M->p_items, pb.items.data(), pb.items.size() * sizeof(PanelItem), cudaMemcpyHostToDevice));
        GCNB_CHECK_CUDA(cudaMalloc(&M->p_subptr, pb.subptr.size() * 4));
        GCNB_CHECK_CUDA(cudaMemcpy(M->p_subptr, pb.subptr.data(), pb.subptr.size() * 4, cudaMemcpyHostToDevice));
        M->bytes += pb.items.size() * sizeof(PanelItem) + pb.subptr.size() * 4;
        if (!pb.split_rows.empty()) {
          GCNB_CHECK_CUDA(cudaMalloc(&M->p_split_rows, pb.split_rows.size() * 4));
          GCNB_CHECK_CUDA(cudaMalloc(&M->p_split_ptr, pb.split_ptr.size() * 4));
          GCNB_CHECK_CUDA(cudaMemcpy(M->p_split_rows, pb.split_rows.data(), pb.split_rows.size() * 4, cudaMemcpyHostToDevice));
          GCNB_CHECK_CUDA(cudaMemcpy(M->p_split_ptr, pb.split_ptr.data(), pb.split_ptr.size() * 4, cudaMemcpyHostToDevice));
          M->bytes += (pb.split_rows.size() + pb.split_ptr.size()) * 4;
        }
      }
    }
  }
  if (!items.empty()) {
    GCNB_CHECK_CUDA(cudaMalloc(&M->items, items.size() * sizeof(WorkItem)));
    GCNB_CHECK_CUDA(cudaMemcpy(M->items, items.data(), items.size() * sizeof(WorkItem),
                               cudaMemcpyHostToDevice));
    M->bytes += items.size() * sizeof(WorkItem);
  }
  if (!split_rows.empty()) {
    GCNB_CHECK_CUDA(cudaMalloc(&M->split_rows, split_rows.size() * 4));
    GCNB_CHECK_CUDA(cudaMalloc(&M->split_ptr, split_ptr.size() * 4));
    GCNB_CHECK_CUDA(cudaMemcpy(M->split_rows, split_rows.data(), split_rows.size() * 4,
                               cudaMemcpyHostToDevice));
    GCNB_CHECK_CUDA(cudaMemcpy(M->split_ptr, split_ptr.data(), split_ptr.size() * 4,
                               cudaMemcpyHostToDevice));
    M->bytes += (split_rows.size() + split_ptr.size()) * 4;
  }
  return GCNB_OK;
}

// CSR transpose by counting sort: columns of every output row come out ascending.
static void transpose_csr(const int32_t* rowptr, const int32_t* col, const float* val, int64_t m,
                          int64_t n, int64_t nnz, std::vector<int32_t>& t_rowptr,
                          std::vector<int32_t>& t_col, std::vector<float>& t_val) {
  t_rowptr.assign((size_t)n + 1, 0);
  t_col.resize((size_t)nnz);
  t_val.resize((size_t)nnz);
  for (int64_t z = 0; z < nnz; ++z) ++t_rowptr[(size_t)col[z] + 1];
  for (int64_t c = 0; c < n; ++c) t_rowptr[c + 1] += t_rowptr[c];
  std::vector<int32_t> cursor(t_rowptr.begin(), t_rowptr.end() - 1);
  for (int64_t r = 0; r < m; ++r)
    for (int64_t z = rowptr[r]; z < rowptr[r + 1]; ++z) {
      const int32_t p = cursor[col[z]]++;
      t_col[p] = (int32_t)r;
      t_val[p] = val[z];
    }
}

// ---- dispatch ------------------------------------------------------------------------------------
struct LaunchArgs {
  const Matrix* M;
  const void* B;
  int64_t ldb;
  float* C;
  int64_t ldc;
  int k;
  const float* bias;
  int epi;
  float* partials;
  cudaStream_t stream;
};

// stream-class launch: ring depth U and CTAs/SM by registers per ring slot (NV*VEC floats + 1)
template <typename TB, int VEC, int NV, int U, int MINB>
static void launch_stream_cfg(const LaunchArgs& a, int64_t gs, unsigned ytiles) {
  const Matrix& M = *a.M;
  // GCNB_L2POL: bit 0 = (col,val) stream and C rows evict_first, bit 1 = feature rows evict_last
  static const int l2pol = getenv("GCNB_L2POL") ? atoi(getenv("GCNB_L2POL")) : 1;
  static const int medium = getenv("GCNB_PACKED_MEDIUM") ? atoi(getenv("GCNB_PACKED_MEDIUM")) : 1;
  const int l2 = ((l2pol & 1) ? kL2Streams : 0) | ((l2pol & 2) ? kL2BLast : 0) | (medium ? 0 : kNoMedium);
  spmm_stream_kernel<TB, VEC, NV, U, MINB><<<dim3((unsigned)gs, ytiles), kStreamWarps * 32, 0, a.stream>>>(
      M.segs, (int)M.n_segs, M.rowptr, M.cv, (const TB*)a.B, (uint32_t)(a.ldb * sizeof(TB)), a.C, a.ldc,
      a.partials, a.k, a.bias, a.epi | l2);
}

template <typename TB, int VEC, int NV>
static int launch_stream(const LaunchArgs& a, int64_t gs, unsigned ytiles) {
  static const int cfg = getenv("GCNB_STREAM_CFG") ? atoi(getenv("GCNB_STREAM_CFG")) : 0;
  constexpr int F = NV * VEC;                        // accumulator registers per lane
  constexpr int R = F * (int)sizeof(TB) / 4;         // registers per ring slot (raw gathered row)
  // measured on the products-shaped graph (profiles/r01_e_stream_cfg_sweep.jsonl, r01_j_*): what
  // counts is resident warps x slots in flight.  fp32 k=100 (F=4,R=4): U=4 at 64 regs / 32 warps per
  // SM beats every other point; bf16 k=100 (F=4,R=2): U=4 at 48 regs / 40 warps.  bf16 with 16-byte
  // loads (F=8,R=4) was 1.6x slower with the U=2 ring the fp32 F=8 case needs, so the ring depth
  // follows R, not F.
  if constexpr (F <= 4) {
    switch (cfg) {
      case 1: launch_stream_cfg<TB, VEC, NV, 4, 10>(a, gs, ytiles); break;   // 48 regs, 40 warps/SM
      case 2: launch_stream_cfg<TB, VEC, NV, 4, 12>(a, gs, ytiles); break;   // 40 regs, 48 warps/SM
      case 3: launch_stream_cfg<TB, VEC, NV, 8, 8>(a, gs, ytiles); break;
      case 4: launch_stream_cfg<TB, VEC, NV, 2, 12>(a, gs, ytiles); break;
      case 5: launch_stream_cfg<TB, VEC, NV, 8, 6>(a, gs, ytiles); break;
      case 6: launch_stream_cfg<TB, VEC, NV, 3, 10>(a, gs, ytiles); break;   // 48 regs, 40 warps/SM, 3 in flight each
      case 7: launch_stream_cfg<TB, VEC, NV, 3, 12>(a, gs, ytiles); break;   // 40 regs, 48 warps/SM
      case 8: launch_stream_cfg<TB, VEC, NV, 3, 8>(a, gs, ytiles); break;
      case 9: launch_stream_cfg<TB, VEC, NV, 2, 16>(a, gs, ytiles); break;   // 32 regs, 64 warps/SM
      case 10: launch_stream_cfg<TB, VEC, NV, 4, 8>(a, gs, ytiles); break;   // 64 regs, 32 warps/SM (the default until r02_au)
      case 11: launch_stream_cfg<TB, VEC, NV, 8, 5>(a, gs, ytiles); break;   // 96 regs, 20 warps/SM, 8 in flight each
      default:
        // fp32: ring 4 at 56 registers / 36 warps per SM since the (col,val) slab is a cp.async ring (no spills any
        // more): products k=100 3.43 mean / 3.20 best against 3.47 / 3.33 at 64 registers / 32 warps
        // (profiles/r02_au_stream_cfg.jsonl; both runs at the board's power cap)
        if constexpr (R <= 2) launch_stream_cfg<TB, VEC, NV, 4, 10>(a, gs, ytiles);
        else launch_stream_cfg<TB, VEC, NV, 4, 9>(a, gs, ytiles);
        break;
    }
  } else if constexpr (F <= 8) {
    switch (cfg) {
      case 1: launch_stream_cfg<TB, VEC, NV, 4, 6>(a, gs, ytiles); break;
      case 2: launch_stream_cfg<TB, VEC, NV, 2, 8>(a, gs, ytiles); break;
      default:
        if constexpr (R <= 4) launch_stream_cfg<TB, VEC, NV, 4, 8>(a, gs, ytiles);   // bf16: 16 ring + 8 acc
        else launch_stream_cfg<TB, VEC, NV, 2, 8>(a, gs, ytiles);
        break;
    }
  } else {
    switch (cfg) {
      case 1: launch_stream_cfg<TB, VEC, NV, 2, 6>(a, gs, ytiles); break;
      case 2: launch_stream_cfg<TB, VEC, NV, 1, 8>(a, gs, ytiles); break;
      default:
        if constexpr (R <= 8) launch_stream_cfg<TB, VEC, NV, 2, 8>(a, gs, ytiles);   // bf16: 16 ring + 16 acc
        else launch_stream_cfg<TB, VEC, NV, 2, 5>(a, gs, ytiles);
        break;
    }
  }
  return GCNB_OK;
}

template <typename TB, int VEC, int G, int NV>
static int launch_one(const LaunchArgs& a) {
  const Matrix& M = *a.M;
  const int W = G * NV * VEC;
  const unsigned ytiles = (unsigned)((a.k + W - 1) / W);
  const int rows_per_cta = kRowWarps * (32 / G);
  const int64_t gx = (M.m + rows_per_cta - 1) / rows_per_cta;
  if (gx > 0x7fffffffLL) return set_error(GCNB_ERR_UNSUPPORTED, "too many rows for one grid");
  const int long_thresh = M.n_items > 0 ? kLongRow : 0x7fffffff;
  if (gx > 0) {
    if constexpr (G == 32) {
      // wide feature rows: warp-per-row kernel (row pitch must fit 32 bits of bytes)
      const int64_t gw = (M.m + kWarpRowWarps - 1) / kWarpRowWarps;
      if (gw > 0x7fffffffLL || a.ldb * (int64_t)sizeof(TB) > 0xffffffffLL)
        return set_error(GCNB_ERR_UNSUPPORTED, "matrix too large for one grid");
      // default 3 = stream class.  (A warp per row with the C load hoisted was tried for the
      // accumulating remote-column block of the row-partitioned SpMM, ~2.6 non-zeros per row:
      // 0.80 ms against the stream kernel's 0.60 ms on 2 GPUs, profiles/r01_m_*.)
      static const int variant = getenv("GCNB_SPMM_VARIANT") ? atoi(getenv("GCNB_SPMM_VARIANT")) : 3;
      if (variant == 3) {
        // stream class covers every row, long ones included: one launch (+ the split-row reduction)
        if (M.n_segs > 0) {
          const int64_t gs = (M.n_segs + kStreamWarps - 1) / kStreamWarps;
          int rc = launch_stream<TB, VEC, NV>(a, gs, ytiles);
          if (rc != GCNB_OK) return rc;
          GCNB_CHECK_CUDA(cudaGetLastError());
          if (M.s_n_split > 0) {
            spmm_reduce_kernel<<<(unsigned)M.s_n_split, 128, 0, a.stream>>>(
                M.s_split_rows, M.s_split_ptr, a.partials, a.C, a.ldc, a.k, a.bias, a.epi);
            GCNB_CHECK_CUDA(cudaGetLastError());
          }
        }
        return GCNB_OK;
      } else
      if (variant == 1)
        spmm_warp_row_kernel<TB, VEC, NV, true><<<dim3((unsigned)gw, ytiles), kWarpRowWarps * 32, 0, a.stream>>>(
            M.rowptr, M.cv, (const TB*)a.B, (uint32_t)(a.ldb * sizeof(TB)), a.C, a.ldc, M.m, a.k, a.bias,
            a.epi, long_thresh);
      else if (variant == 0)
        spmm_warp_row_kernel<TB, VEC, NV, false><<<dim3((unsigned)gw, ytiles), kWarpRowWarps * 32, 0, a.stream>>>(
            M.rowptr, M.cv, (const TB*)a.B, (uint32_t)(a.ldb * sizeof(TB)), a.C, a.ldc, M.m, a.k, a.bias,
            a.epi, long_thresh);
      else
        spmm_rows_kernel<TB, VEC, G, NV><<<dim3((unsigned)gx, ytiles), kRowWarps * 32, 0, a.stream>>>(
            M.rowptr, M.cv, (const TB*)a.B, a.ldb, a.C, a.ldc, M.m, a.k, a.bias, a.epi, long_thresh);
    } else {
      spmm_rows_kernel<TB, VEC, G, NV><<<dim3((unsigned)gx, ytiles), kRowWarps * 32, 0, a.stream>>>(
          M.rowptr, M.cv, (const TB*)a.B, a.ldb, a.C, a.ldc, M.m, a.k, a.bias, a.epi, long_thresh);
    }
    GCNB_CHECK_CUDA(cudaGetLastError());
  }
  if (M.n_items > 0) {
    spmm_block_kernel<TB, VEC, G, NV>
        <<<dim3((unsigned)M.n_items, ytiles), kBlockWarps * 32, 0, a.stream>>>(
            M.items, M.cv, (const TB*)a.B, a.ldb, a.C, a.ldc, a.partials, a.k, a.bias, a.epi);
    GCNB_CHECK_CUDA(cudaGetLastError());
    if (M.n_split > 0) {
      spmm_reduce_kernel<<<(unsigned)M.n_split, 128, 0, a.stream>>>(
          M.split_rows, M.split_ptr, a.partials, a.C, a.ldc, a.k, a.bias, a.epi);
      GCNB_CHECK_CUDA(cudaGetLastError());
    }
  }
  return GCNB_OK;
}

// narrow feature rows: S non-zeros per warp-wide gather (spmm_stream_packed_kernel), one launch for
// every row class (+ the split-row reduction)
template <typename TB, int VEC, int S>
static int launch_packed(const LaunchArgs& a) {
  const Matrix& M = *a.M;
  if (M.n_segs == 0) return GCNB_OK;
  constexpr int U = 4;
  constexpr int MINB = (sizeof(TB) == 2 && VEC <= 4) ? 10 : 8;
  const int64_t gs = (M.n_segs + kStreamWarps - 1) / kStreamWarps;
  if (gs > 0x7fffffffLL || a.ldb * (int64_t)sizeof(TB) > 0xffffffffLL)
    return set_error(GCNB_ERR_UNSUPPORTED, "matrix too large for one grid");
  static const int l2pol = getenv("GCNB_L2POL") ? atoi(getenv("GCNB_L2POL")) : 1;
  static const int medium = getenv("GCNB_PACKED_MEDIUM") ? atoi(getenv("GCNB_PACKED_MEDIUM")) : 1;
  const int l2 = ((l2pol & 1) ? kL2Streams : 0) | ((l2pol & 2) ? kL2BLast : 0) | (medium ? 0 : kNoMedium);
  if constexpr (sizeof(TB) == 4 && VEC == 8) {
    // 256-bit loads: a ring slot is 8 registers.  GCNB_PACKED_CFG picks ring depth / CTAs per SM (experiments)
    static const int pcfg = getenv("GCNB_PACKED_CFG") ? atoi(getenv("GCNB_PACKED_CFG")) : 0;
    auto go = [&](auto kern) {
      kern<<<(unsigned)gs, kStreamWarps * 32, 0, a.stream>>>(
          M.segs, (int)M.n_segs, M.rowptr, M.cv, (const TB*)a.B, (uint32_t)(a.ldb * sizeof(TB)), a.C, a.ldc,
          a.partials, a.k, a.bias, a.epi | l2);
    };
    // registers (ptxas, S = 2): U=2 at 6 CTAs/SM 78, U=4 at 4 CTAs/SM 118, U=8 at 3 CTAs/SM 162, none spilling;
    // U=2 at 8 CTAs/SM (64 registers) spills a ring slot
    constexpr int U8 = 8 * S <= 32 ? 8 : 32 / S;   // a ring turn must fit a 32-entry slab
    if (pcfg == 1) go(spmm_stream_packed_kernel<TB, VEC, S, 4, 4>);
    else if (pcfg == 2) go(spmm_stream_packed_kernel<TB, VEC, S, U8, 3>);
    else if (pcfg == 3) go(spmm_stream_packed_kernel<TB, VEC, S, 2, 8>);
    else if (pcfg == 4) go(spmm_stream_packed_kernel<TB, VEC, S, 4, 5>);
    else go(spmm_stream_packed_kernel<TB, VEC, S, 2, 6>);
  } else {
    spmm_stream_packed_kernel<TB, VEC, S, U, MINB><<<(unsigned)gs, kStreamWarps * 32, 0, a.stream>>>(
        M.segs, (int)M.n_segs, M.rowptr, M.cv, (const TB*)a.B, (uint32_t)(a.ldb * sizeof(TB)), a.C, a.ldc,
        a.partials, a.k, a.bias, a.epi | l2);
  }
  GCNB_CHECK_CUDA(cudaGetLastError());
  if (M.s_n_split > 0) {
    spmm_reduce_kernel<<<(unsigned)M.s_n_split, 128, 0, a.stream>>>(
        M.s_split_rows, M.s_split_ptr, a.partials, a.C, a.ldc, a.k, a.bias, a.epi);
    GCNB_CHECK_CUDA(cudaGetLastError());
  }
  return GCNB_OK;
}

// panel class: one CTA per item, 32 warps = the 32 rows of the panel, column tiles of 32 vectors in grid.y
template <typename TB, int VEC>
static int launch_panel(const LaunchArgs& a) {
  const Matrix& M = *a.M;
  static const int l2pol = getenv("GCNB_L2POL") ? atoi(getenv("GCNB_L2POL")) : 1;
  const int l2 = ((l2pol & 1) ? kL2Streams : 0) | ((l2pol & 2) ? kL2BLast : 0);
  const unsigned ytiles = (unsigned)((a.k + 32 * VEC - 1) / (32 * VEC));
  if (a.ldb * (int64_t)sizeof(TB) > 0xffffffffLL || ytiles > 65535)
    return set_error(GCNB_ERR_UNSUPPORTED, "matrix too large for one grid");
  spmm_panel_kernel<TB, VEC, 4><<<dim3((unsigned)M.p_n_items, ytiles), kPanelRows * 32, 0, a.stream>>>(
      M.p_items, M.p_subptr, M.cv, (const TB*)a.B, (uint32_t)(a.ldb * sizeof(TB)), a.C, a.ldc, a.partials,
      (int)M.m, a.k, a.bias, a.epi | l2);
  GCNB_CHECK_CUDA(cudaGetLastError());
  if (M.p_n_split > 0) {
    spmm_reduce_kernel<<<(unsigned)M.p_n_split, 128, 0, a.stream>>>(
        M.p_split_rows, M.p_split_ptr, a.partials, a.C, a.ldc, a.k, a.bias, a.epi);
    GCNB_CHECK_CUDA(cudaGetLastError());
  }
  return GCNB_OK;
}

template <typename TB, int VEC, int NVMAX>
static int launch_vec(const LaunchArgs& a) {
  const int nvec = (a.k + VEC - 1) / VEC;  // vectors per feature row
  // wide feature rows on a matrix whose panels share columns: panel class (not in accumulate mode: the
  // remote-column block of the partitioned SpMM has a handful of non-zeros per row)
  if (a.M->p_n_items > 0 && nvec > 16 && !(a.epi & GCNB_EPI_ACCUM)) return launch_panel<TB, VEC>(a);
  // GCNB_PACKED = largest S the packed stream kernel may use (0: lane-group / block kernels instead)
  static const int packed = getenv("GCNB_PACKED") ? atoi(getenv("GCNB_PACKED")) : 8;
  // Measured (profiles/r01_v_*): products-shaped (50 nnz/row) k=32: 1.86 ms packed vs 2.87 ms lane-group
  // kernel, k=48: 2.58 vs 3.73; Reddit-shaped k=48: 1.05 vs 1.53; but arxiv-shaped (15 nnz/row) k=40:
  // 0.097 vs 0.080 — with short rows the fold at every row end outweighs the saved requests, and the
  // lane-group kernel (several rows side by side in a warp) wins.
  if (packed >= 2 && nvec <= 16 && a.M->nnz >= 24 * a.M->m) {
    if (nvec <= 4 && packed >= 8) return launch_packed<TB, VEC, 8>(a);
    if (nvec <= 8 && packed >= 4) return launch_packed<TB, VEC, 4>(a);
    return launch_packed<TB, VEC, 2>(a);
  }
  // rows needing more than `stream_min` lanes go to the whole-warp stream class, narrower ones to the
  // lane-group kernel (several rows per warp)
  static const int stream_min = getenv("GCNB_STREAM_MIN_NVEC") ? atoi(getenv("GCNB_STREAM_MIN_NVEC")) : 16;
  if (nvec <= 4 && stream_min >= 4) return launch_one<TB, VEC, 4, 1>(a);
  if (nvec <= 8 && stream_min >= 8) return launch_one<TB, VEC, 8, 1>(a);
  if (nvec <= 16 && stream_min >= 16) return launch_one<TB, VEC, 16, 1>(a);
  if (nvec <= 32) return launch_one<TB, VEC, 32, 1>(a);
  // wide rows: whole warp per row, NV vectors per lane, as few column tiles as possible.  GCNB_WIDE_NV caps NV
  // (more, narrower column slices in grid.y, which run one after the other): tried for the Reddit-shaped graph
  // at k = 602, whose operand (561 MB) misses L2 (71 GB of DRAM reads per launch with two 384-column slices,
  // profiles/r02_d_ncu_reddit_deg_k602.txt) — five 128-column slices were SLOWER (22.3 vs 20.7 ms,
  // profiles/r02_e_reddit_k602_*): a 119 MB slice does not stay in the 126 MB L2 either and the (col,val)
  // stream is re-read per slice.  So the cap stays an experiment switch (0 = none).
  static const int wide_nv = getenv("GCNB_WIDE_NV") ? atoi(getenv("GCNB_WIDE_NV")) : 0;
  int nvmax = NVMAX;
  if (wide_nv > 0) nvmax = std::min(NVMAX, wide_nv);
  const int tiles = (nvec + 32 * nvmax - 1) / (32 * nvmax);
  const int nv = (nvec + 32 * tiles - 1) / (32 * tiles);
  if (nv <= 1) return launch_one<TB, VEC, 32, 1>(a);
  if (nv == 2) return launch_one<TB, VEC, 32, 2>(a);
  if constexpr (NVMAX >= 4) {
    if (nv == 3) return launch_one<TB, VEC, 32, 3>(a);
    return launch_one<TB, VEC, 32, 4>(a);
  } else {
    return launch_one<TB, VEC, 32, 2>(a);
  }
}

static inline bool aligned(const void* p, size_t a) { return ((uintptr_t)p % a) == 0; }

// fp32 feature rows through 256-bit loads (VEC = 8): half the warp-wide gather requests of the 128-bit kernels.
// Needs 32-byte aligned rows (B and its pitch) whose pitch covers k rounded up to 8.  Only the stream classes:
//   k <= 32 / 64 / 128 on matrices with >= 24 nnz/row: packed, 8 / 4 / 2 rows per request
//   k <= 256: one request per row (NV = 1), wider: NV = 2 per column tile of 512
// returns 1 when the shape is left to the 128-bit kernels
static int launch_vec8(const LaunchArgs& a, int mask) {
  const int nvec = (a.k + 7) / 8;
  const Matrix& M = *a.M;
  if (M.n_segs == 0) return 1;
  if (M.p_n_items > 0 && !(a.epi & GCNB_EPI_ACCUM)) return 1;   // panel class chosen at plan time
  static const int packed = getenv("GCNB_PACKED") ? atoi(getenv("GCNB_PACKED")) : 8;
  if (nvec <= 16) {
    if (!(mask & 1) || packed < 2 || M.nnz < 24 * M.m) return 1;
    if (nvec <= 4 && packed >= 8) return launch_packed<float, 8, 8>(a);
    if (nvec <= 8 && packed >= 4) return launch_packed<float, 8, 4>(a);
    return launch_packed<float, 8, 2>(a);
  }
  if (!(mask & 2)) return 1;
  if (a.ldb * 4 > 0xffffffffLL) return 1;
  const int tiles = (nvec + 63) / 64;
  const int nv = (nvec + 32 * tiles - 1) / (32 * tiles);
  const int64_t gs = (M.n_segs + kStreamWarps - 1) / kStreamWarps;
  if (gs > 0x7fffffffLL) return 1;
  const unsigned ytiles = (unsigned)((a.k + 32 * nv * 8 - 1) / (32 * nv * 8));
  int rc = nv <= 1 ? launch_stream<float, 8, 1>(a, gs, ytiles) : launch_stream<float, 8, 2>(a, gs, ytiles);
  if (rc != GCNB_OK) return rc;
  GCNB_CHECK_CUDA(cudaGetLastError());
  if (M.s_n_split > 0) {
    spmm_reduce_kernel<<<(unsigned)M.s_n_split, 128, 0, a.stream>>>(
        M.s_split_rows, M.s_split_ptr, a.partials, a.C, a.ldc, a.k, a.bias, a.epi);
    GCNB_CHECK_CUDA(cudaGetLastError());
  }
  return GCNB_OK;
}

static int spmm_dispatch(const Matrix& M, const void* B, int b_dtype, int64_t ldb, float* C,
                         int64_t ldc, int k, const float* bias, int epi, cudaStream_t stream) {
  if (M.m == 0 || k == 0) return GCNB_OK;
  float* partials = nullptr;
  const int64_t n_part = std::max(std::max(M.n_partials, M.s_n_partials), M.p_n_partials);
  if (n_part > 0)
    GCNB_CHECK_CUDA(cudaMallocAsync(&partials, (size_t)n_part * k * sizeof(float), stream));
  LaunchArgs a{&M, B, ldb, C, ldc, k, bias, epi, partials, stream};
  int rc;
  // vector width: limited by k, both leading dimensions and base alignment (C is stored with
  // min(VEC,4) floats per store)
  static const int max_vec = getenv("GCNB_MAX_VEC") ? atoi(getenv("GCNB_MAX_VEC")) : 4;   // experiments
  // GCNB_F32_V8: bit 0 = 256-bit packed kernels (k <= 128), bit 1 = 256-bit stream kernel (k > 128)
  static const int f32_v8 = getenv("GCNB_F32_V8") ? atoi(getenv("GCNB_F32_V8")) : 0;
  if (b_dtype == GCNB_F32 && f32_v8 && k % 4 == 0 && ldb % 8 == 0 && ldb >= (k + 7) / 8 * 8 && ldc % 4 == 0 &&
      aligned(B, 32) && aligned(C, 16)) {
    rc = launch_vec8(a, f32_v8);
    if (rc != 1) {
      if (partials) cudaFreeAsync(partials, stream);
      return rc;
    }
  }
  if (b_dtype == GCNB_F32) {
    if (max_vec >= 4 && k % 4 == 0 && ldb % 4 == 0 && ldc % 4 == 0 && aligned(B, 16) && aligned(C, 16))
      rc = launch_vec<float, 4, 4>(a);
    else if (k % 2 == 0 && ldb % 2 == 0 && ldc % 2 == 0 && aligned(B, 8) && aligned(C, 8))
      rc = launch_vec<float, 2, 4>(a);
    else
      rc = launch_vec<float, 1, 4>(a);
  } else {
    // wide rows (stream class): 8-byte loads over up to 32 lanes measured 1.6x faster than 16-byte
    // loads over half as many lanes (products-shaped, k = 128: 3.1 vs 4.9 ms, profiles/r01_j_*), so
    // 16-byte loads are kept for the narrow lane-group kernel only
    const bool v4ok = k % 4 == 0 && ldb % 4 == 0 && ldc % 4 == 0 && aligned(B, 8) && aligned(C, 16);
    static const int bf16_v8_max = getenv("GCNB_BF16_V8_MAX") ? atoi(getenv("GCNB_BF16_V8_MAX")) : 64;
    if (v4ok && k > bf16_v8_max)
      rc = launch_vec<__nv_bfloat16, 4, 4>(a);
    else if (k % 8 == 0 && ldb % 8 == 0 && ldc % 4 == 0 && aligned(B, 16) && aligned(C, 16))
      rc = launch_vec<__nv_bfloat16, 8, 2>(a);
    else if (k % 4 == 0 && ldb % 4 == 0 && ldc % 4 == 0 && aligned(B, 8) && aligned(C, 16))
      rc = launch_vec<__nv_bfloat16, 4, 4>(a);
    else if (k % 2 == 0 && ldb % 2 == 0 && ldc % 2 == 0 && aligned(B, 4) && aligned(C, 8))
      rc = launch_vec<__nv_bfloat16, 2, 4>(a);
    else
      rc = launch_vec<__nv_bfloat16, 1, 4>(a);
  }
  if (partials) cudaFreeAsync(partials, stream);
  return rc;
}

static int check_operands(const gcnb_plan_t* plan, const void* B, int b_dtype, int64_t ldb,
                          const float* C, int64_t ldc, int k, const float* bias, int epi) {
  GCNB_REQUIRE(plan != nullptr, "plan is null");
  GCNB_REQUIRE(k >= 0, "k = %d is negative", k);
  GCNB_REQUIRE(b_dtype == GCNB_F32 || b_dtype == GCNB_BF16, "unknown b_dtype %d", b_dtype);
  GCNB_REQUIRE(ldb >= k && ldc >= k, "leading dimensions (%lld, %lld) smaller than k = %d",
               (long long)ldb, (long long)ldc, k);
  GCNB_REQUIRE(k == 0 || (B != nullptr && C != nullptr), "B or C is null");
  GCNB_REQUIRE(!(epi & GCNB_EPI_BIAS) || bias != nullptr, "GCNB_EPI_BIAS without a bias pointer");
  GCNB_REQUIRE((epi & ~(GCNB_EPI_BIAS | GCNB_EPI_RELU | GCNB_EPI_ACCUM)) == 0, "unknown epilogue flags %d", epi);
  return GCNB_OK;
}

}  // namespace gcnb

using namespace gcnb;

extern "C" {

int gcnb_plan_create(const int32_t* rowptr, const int32_t* col, const float* val, int64_t m,
                     int64_t n, int64_t nnz, int flags, int device, gcnb_plan_t** plan_out) {
  GCNB_REQUIRE(plan_out != nullptr, "plan_out is null");
  *plan_out = nullptr;
  GCNB_REQUIRE(m >= 0 && n >= 0 && nnz >= 0, "negative size (m=%lld n=%lld nnz=%lld)",
               (long long)m, (long long)n, (long long)nnz);
  GCNB_REQUIRE(rowptr != nullptr, "rowptr is null");
  GCNB_REQUIRE(nnz == 0 || (col != nullptr && val != nullptr), "col or val is null");
  GCNB_REQUIRE(m < 0x7fffffffLL && n < 0x7fffffffLL && nnz < 0x7ffffff0LL,
               "int32 CSR limits exceeded (m=%lld n=%lld nnz=%lld)", (long long)m, (long long)n,
               (long long)nnz);
  GCNB_REQUIRE(rowptr[0] == 0 && rowptr[m] == nnz, "rowptr[0]=%d rowptr[m]=%d, expected 0 and %lld",
               rowptr[0], rowptr[m], (long long)nnz);
  for (int64_t r = 0; r < m; ++r) {
    GCNB_REQUIRE(rowptr[r + 1] >= rowptr[r], "rowptr decreases at row %lld", (long long)r);
    for (int64_t z = rowptr[r]; z < rowptr[r + 1]; ++z) {
      GCNB_REQUIRE(col[z] >= 0 && col[z] < n, "column %d out of range at nz %lld", col[z],
                   (long long)z);
      GCNB_REQUIRE(z == rowptr[r] || col[z] > col[z - 1],
                   "columns of row %lld are not strictly ascending", (long long)r);
    }
  }
  int rc = require_gpu();
  if (rc != GCNB_OK) return rc;
  int ndev = gcnb_device_count();
  GCNB_REQUIRE(device >= 0 && device < ndev, "device %d out of range (%d visible)", device, ndev);
  DeviceGuard guard(device);
  GCNB_CHECK_CUDA(guard.status);

  gcnb_plan* p = new (std::nothrow) gcnb_plan();
  if (!p) return set_error(GCNB_ERR_ALLOC, "out of host memory");
  p->device = device;
  p->n_sm = sm_count(device);
  rc = upload_matrix(rowptr, col, val, m, n, nnz, p->n_sm, flags, &p->a);
  if (rc != GCNB_OK) { gcnb_plan_destroy(p); return rc; }

  if (flags & GCNB_PLAN_ASSUME_SYMMETRIC) {
    if (m != n) { gcnb_plan_destroy(p); return set_error(GCNB_ERR_ARG, "ASSUME_SYMMETRIC with m != n"); }
    p->symmetric = true;
  } else if (!(flags & GCNB_PLAN_NO_TRANSPOSE)) {
    std::vector<int32_t> t_rowptr, t_col;
    std::vector<float> t_val;
    transpose_csr(rowptr, col, val, m, n, nnz, t_rowptr, t_col, t_val);
    if (m == n && memcmp(t_rowptr.data(), rowptr, (size_t)(m + 1) * 4) == 0 &&
        (nnz == 0 || (memcmp(t_col.data(), col, (size_t)nnz * 4) == 0 &&
                      memcmp(t_val.data(), val, (size_t)nnz * 4) == 0))) {
      p->symmetric = true;  // bit-identical transpose: A^T·dY can run on A
    } else {
      rc = upload_matrix(t_rowptr.data(), t_col.data(), t_val.data(), n, m, nnz, p->n_sm, flags, &p->at);
      if (rc != GCNB_OK) { gcnb_plan_destroy(p); return rc; }
      p->has_t = true;
    }
  }
  *plan_out = p;
  return GCNB_OK;
}

int gcnb_plan_destroy(gcnb_plan_t* plan) {
  if (!plan) return GCNB_OK;
  DeviceGuard guard(plan->device);
  plan->a.release();
  plan->at.release();
  cudaFree(plan->stage_b);
  cudaFree(plan->stage_c);
  delete plan;
  return GCNB_OK;
}

int gcnb_plan_info(const gcnb_plan_t* plan, int64_t info[16]) {
  GCNB_REQUIRE(plan != nullptr && info != nullptr, "null argument");
  for (int i = 0; i < 16; ++i) info[i] = 0;
  info[0] = plan->a.m; info[1] = plan->a.n; info[2] = plan->a.nnz;
  info[3] = plan->symmetric; info[4] = plan->has_t;
  info[5] = plan->a.m - plan->a.n_long_rows; info[6] = plan->a.n_long_rows;
  info[7] = plan->a.n_items; info[8] = plan->a.max_row_nnz;
  info[9] = (int64_t)(plan->a.bytes + plan->at.bytes + plan->stage_b_bytes + plan->stage_c_bytes);
  info[10] = plan->n_sm; info[11] = plan->a.n_split;
  info[12] = plan->a.n_segs; info[13] = plan->a.s_n_split;
  info[14] = plan->a.p_n_items; info[15] = plan->a.p_distinct;
  return GCNB_OK;
}

int gcnb_spmm(const gcnb_plan_t* plan, const void* B, int b_dtype, int64_t ldb, float* C,
              int64_t ldc, int k, const float* bias, int epilogue, void* stream) {
  int rc = check_operands(plan, B, b_dtype, ldb, C, ldc, k, bias, epilogue);
  if (rc != GCNB_OK) return rc;
  DeviceGuard guard(plan->device);
  GCNB_CHECK_CUDA(guard.status);
  return spmm_dispatch(plan->a, B, b_dtype, ldb, C, ldc, k, bias, epilogue, (cudaStream_t)stream);
}

int gcnb_spmm_t(const gcnb_plan_t* plan, const void* B, int b_dtype, int64_t ldb, float* C,
                int64_t ldc, int k, void* stream) {
  int rc = check_operands(plan, B, b_dtype, ldb, C, ldc, k, nullptr, 0);
  if (rc != GCNB_OK) return rc;
  if (!plan->symmetric && !plan->has_t)
    return set_error(GCNB_ERR_ARG, "plan was created with GCNB_PLAN_NO_TRANSPOSE");
  DeviceGuard guard(plan->device);
  GCNB_CHECK_CUDA(guard.status);
  const Matrix& M = plan->symmetric ? plan->a : plan->at;
  return spmm_dispatch(M, B, b_dtype, ldb, C, ldc, k, nullptr, 0, (cudaStream_t)stream);
}

int gcnb_spmm_host(gcnb_plan_t* plan, const float* B_host, float* C_host, int k, int transpose,
                   void* stream) {
  GCNB_REQUIRE(plan != nullptr, "plan is null");
  GCNB_REQUIRE(k >= 0, "k = %d is negative", k);
  GCNB_REQUIRE(k == 0 || (B_host && C_host), "host buffer is null");
  if (transpose && !plan->symmetric && !plan->has_t)
    return set_error(GCNB_ERR_ARG, "plan was created with GCNB_PLAN_NO_TRANSPOSE");
  DeviceGuard guard(plan->device);
  GCNB_CHECK_CUDA(guard.status);
  const Matrix& M = (transpose && !plan->symmetric) ? plan->at : plan->a;
  if (k == 0 || M.m == 0) return GCNB_OK;
  cudaStream_t s = (cudaStream_t)stream;
  const size_t bb = (size_t)M.n * k * sizeof(float), cb = (size_t)M.m * k * sizeof(float);
  if (bb > plan->stage_b_bytes) {
    cudaFree(plan->stage_b);
    plan->stage_b = nullptr; plan->stage_b_bytes = 0;
    GCNB_CHECK_CUDA(cudaMalloc(&plan->stage_b, bb));
    plan->stage_b_bytes = bb;
  }
  if (cb > plan->stage_c_bytes) {
    cudaFree(plan->stage_c);
    plan->stage_c = nullptr; plan->stage_c_bytes = 0;
    GCNB_CHECK_CUDA(cudaMalloc(&plan->stage_c, cb));
    plan->stage_c_bytes = cb;
  }
  if (bb) GCNB_CHECK_CUDA(cudaMemcpyAsync(plan->stage_b, B_host, bb, cudaMemcpyHostToDevice, s));
  int rc = spmm_dispatch(M, plan->stage_b, GCNB_F32, k, plan->stage_c, k, k, nullptr, 0, s);
  if (rc != GCNB_OK) return rc;
  GCNB_CHECK_CUDA(cudaMemcpyAsync(C_host, plan->stage_c, cb, cudaMemcpyDeviceToHost, s));
  GCNB_CHECK_CUDA(cudaStreamSynchronize(s));
  return GCNB_OK;
}

int gcnb_permute_rows(const float* X, const int32_t* vomp, float* out, int64_t n, int k,
                      void* stream) {
  GCNB_REQUIRE(n >= 0 && k >= 0, "negative size");
  if (n == 0 || k == 0) return GCNB_OK;
  GCNB_REQUIRE(X && vomp && out, "null pointer");
  GCNB_REQUIRE(X != out, "gcnb_permute_rows is out of place");
  int rc = require_gpu();
  if (rc != GCNB_OK) return rc;
  int dev = pointer_device(X);
  GCNB_REQUIRE(dev >= 0, "X is not a device pointer");
  DeviceGuard guard(dev);
  GCNB_CHECK_CUDA(guard.status);
  const int vec4 = (k % 4 == 0) && aligned(X, 16) && aligned(out, 16);
  const int64_t blocks = (n + 7) / 8;
  permute_rows_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(X, vomp, out, n, k, vec4);
  GCNB_CHECK_CUDA(cudaGetLastError());
  return GCNB_OK;
}

}  // extern "C"
