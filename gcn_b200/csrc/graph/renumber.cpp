// CSR renumbering: apply a vertex permutation (renumber.cu:233-318 perm_apply, and the
// identical tails of gorder() :190-229 and dfs() :97-154) and the DFS pre-order
// (renumber.cu:56-95).
#include <algorithm>
#include <memory>
#include <utility>

#include "graph.h"

namespace gcnb {
namespace graph {

void perm_apply(int32_t* rowptr, int32_t* col, float* val, const int32_t* vomp, int64_t m,
                int64_t nnz) {
  // new row i is old row vomp[i]; every column id goes through old->new and each row is
  // re-sorted by its new column ids (the tile builder needs ascending columns).
  std::vector<int32_t> to_new((size_t)m);
  std::vector<int32_t> new_rowptr((size_t)m + 1);
  new_rowptr[0] = 0;
  for (int64_t i = 0; i < m; ++i) {
    const int32_t old = vomp[i];
    to_new[(size_t)old] = (int32_t)i;
    new_rowptr[i + 1] = new_rowptr[i] + (rowptr[old + 1] - rowptr[old]);
  }
  // uninitialised scratch (every element is written below): no 8·nnz-byte memset
  std::unique_ptr<int32_t[]> new_col(new int32_t[(size_t)std::max<int64_t>(nnz, 1)]);
  std::unique_ptr<float[]> new_val(new float[(size_t)std::max<int64_t>(nnz, 1)]);
  // rows are independent: one contiguous range of new rows per host thread, balanced by non-zeros
  parallel_rows(m, new_rowptr, [&](int64_t r0, int64_t r1) {
    std::vector<std::pair<int32_t, float>> row;
    for (int64_t i = r0; i < r1; ++i) {
      const int32_t old = vomp[i];
      const int64_t b = rowptr[old], e = rowptr[old + 1];
      row.clear();
      for (int64_t z = b; z < e; ++z) row.emplace_back(to_new[(size_t)col[z]], val[z]);
      // column ids within a row are distinct, so the order is fully determined
      std::sort(row.begin(), row.end(),
                [](const std::pair<int32_t, float>& x, const std::pair<int32_t, float>& y) {
                  return x.first < y.first;
                });
      int64_t p = new_rowptr[i];
      for (const auto& cv : row) {
        new_col[(size_t)p] = cv.first;
        new_val[(size_t)p] = cv.second;
        ++p;
      }
    }
  });
  std::copy(new_rowptr.begin(), new_rowptr.end(), rowptr);
  parallel_rows(m, new_rowptr, [&](int64_t r0, int64_t r1) {
    const int64_t b = new_rowptr[r0], e = new_rowptr[r1];
    std::copy(new_col.get() + b, new_col.get() + e, col + b);
    std::copy(new_val.get() + b, new_val.get() + e, val + b);
  });
}

void dfs_order(const CsrView& g, int32_t* vomp) {
  const int64_t n = g.m;
  std::vector<uint8_t> seen((size_t)n, 0);
  std::vector<std::pair<int32_t, int32_t>> stack;  // (vertex, next edge to look at)
  int64_t placed = 0;
  for (int64_t root = 0; root < n; ++root) {
    if (seen[root]) continue;
    seen[root] = 1;
    vomp[placed++] = (int32_t)root;
    stack.emplace_back((int32_t)root, g.rowptr[root]);
    while (!stack.empty()) {
      auto& top = stack.back();
      const int32_t end = g.rowptr[top.first + 1];
      int32_t z = top.second;
      while (z < end && seen[(size_t)g.col[z]]) ++z;
      if (z >= end) {
        stack.pop_back();
        continue;
      }
      const int32_t v = g.col[z];
      top.second = z + 1;
      seen[(size_t)v] = 1;
      vomp[placed++] = v;
      stack.emplace_back(v, g.rowptr[v]);
    }
  }
}

}  // namespace graph
}  // namespace gcnb
