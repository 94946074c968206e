// Reference tile format: CSR -> column-ordered segments of tm-row panels (tile.cu:11-103) and
// the per-SM segment buckets (tile.cu:118-161).  The engine itself does not consume this format
// (its layout is the plan, spmm.cu); it is produced for parity and for the compat shim.
//
// The reference collects the panel's occupied columns in a std::map and stages every row's pairs
// in per-row vectors.  Here a panel is a tm-way merge over the rows' CSR cursors: because each
// row is consumed in column order, the part of row i that belongs to the open segment is simply
// the CSR range [seg_start[i], cursor[i]), so nothing is staged.
#include <algorithm>

#include "graph.h"

namespace gcnb {
namespace graph {

static int bit_width_u32(uint32_t x) {
  int w = 0;
  while (x) { ++w; x >>= 1; }
  return w;
}

int64_t build_segments(const int32_t* rowptr, const int32_t* col, const float* val, int64_t m,
                       int tm, int64_t first_panel, int64_t last_panel, int64_t start_seg,
                       bool cumulative, const TileOut* out) {
  constexpr int kNnzLimit = 128;                       // tile.cu:23
  const int kSlack = (int)(0.1 * kNnzLimit);           // tile.cu:24  -> 12
  const int32_t pad_row = (int32_t)(1u << (bit_width_u32((uint32_t)m) + 1));  // tile.cu:95
  int64_t seg = start_seg;
  // seg_rowPtr value the next segment starts from: the last entry of the previous segment
  // (tile.cu:56-58)
  int64_t begin = 0;
  if (seg > 0 && out && out->seg_rowPtr) begin = out->seg_rowPtr[(int64_t)(tm + 1) * seg - 1];
  std::vector<int64_t> cursor((size_t)tm), seg_start((size_t)tm), row_end((size_t)tm);

  for (int64_t panel = first_panel; panel < last_panel; ++panel) {
    const int64_t r0 = panel * tm, r1 = std::min(m, r0 + tm);
    const int rows = (int)(r1 - r0);
    if (rowptr[r1] == rowptr[r0]) continue;  // empty panel: undefined in the reference (tile.cu:35)
    int32_t last_col = -1;
    for (int i = 0; i < rows; ++i) {
      cursor[i] = seg_start[i] = rowptr[r0 + i];
      row_end[i] = rowptr[r0 + i + 1];
      if (row_end[i] > cursor[i]) last_col = std::max(last_col, col[row_end[i] - 1]);
    }
    int in_seg = 0;
    for (;;) {
      int32_t j = INT32_MAX;
      for (int i = 0; i < rows; ++i)
        if (cursor[i] < row_end[i]) j = std::min(j, col[cursor[i]]);
      if (j == INT32_MAX) break;
      for (int i = 0; i < rows; ++i)
        if (cursor[i] < row_end[i] && col[cursor[i]] == j) { ++cursor[i]; ++in_seg; }
      const bool close = (j == last_col && in_seg) || (kNnzLimit - in_seg) <= kSlack ||
                         in_seg > kNnzLimit;  // tile.cu:51
      if (!close) continue;
      if (out && out->seg_rowPtr) {
        int32_t* srp = out->seg_rowPtr + (int64_t)(tm + 1) * seg;
        srp[0] = (int32_t)begin;
        int64_t running = 0;
        for (int i = 0; i < tm; ++i) {
          int64_t len = 0;
          if (i < rows) {
            len = cursor[i] - seg_start[i];
            for (int64_t z = seg_start[i]; z < cursor[i]; ++z) {
              out->segNzCV[(begin + running) * 2] = (float)col[z];  // column id as float, tile.cu:67
              out->segNzCV[(begin + running) * 2 + 1] = val[z];
              ++running;
            }
          }
          // corrected: running count (what flexspmm.cu:63-64 and tile.cu:138 index with);
          // as shipped (tile.cu:71): begin + this row's count only
          srp[i + 1] = (int32_t)(begin + (cumulative ? running : len));
        }
        begin = srp[tm];
      } else {
        begin += in_seg;
      }
      if (out && out->segVoMap) {
        int32_t* vm = out->segVoMap + (int64_t)tm * seg;
        for (int i = 0; i < tm; ++i) {
          if (i < rows) {
            const int64_t got = cursor[i] - seg_start[i];
            const int64_t full = row_end[i] - rowptr[r0 + i];
            const int32_t r = (int32_t)(r0 + i);
            // a row only partly inside this segment must be accumulated: MSB flag (tile.cu:79-86)
            vm[i] = (got < full) ? (int32_t)((uint32_t)r | 0x80000000u) : r;
          } else {
            vm[i] = pad_row;
          }
        }
      }
      for (int i = 0; i < rows; ++i) seg_start[i] = cursor[i];
      in_seg = 0;
      ++seg;
    }
  }
  return seg;
}

void build_buckets(const int32_t* seg_rowPtr, int64_t n_segs, int64_t nnz, int tm, int n_sm,
                   int32_t* grouped_tailSeg, int32_t* next_seg) {
  // tile.cu:128-161: n_sm contiguous buckets of about 0.95 * nnz / n_sm non-zeros each, plus a
  // last bucket holding whatever is left for SMs that finish early.
  const int64_t workload = nnz / n_sm;
  const int64_t target = (int64_t)(0.95 * (double)workload);
  auto seg_nnz = [&](int64_t s) -> int64_t {
    if (s >= n_segs) return 0;
    return (int64_t)seg_rowPtr[(s + 1) * (tm + 1) - 1] - seg_rowPtr[s * (tm + 1)];
  };
  int64_t head = 0;
  for (int i = 0; i < n_sm; ++i) {
    next_seg[i] = (int32_t)head;
    int64_t nz = seg_nnz(head);
    int64_t tail = head + 1;
    while (tail < n_segs && nz < target) nz += seg_nnz(tail++);
    grouped_tailSeg[i] = (int32_t)std::min(n_segs, tail);
    head = tail;
  }
  next_seg[n_sm] = (int32_t)head;
  grouped_tailSeg[n_sm] = (int32_t)n_segs;
}

}  // namespace graph
}  // namespace gcnb
