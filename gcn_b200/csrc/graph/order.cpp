// Degree and reverse-Cuthill-McKee vertex orders, bit-exact with the reference's
// order_deg.cu:8-45, order_rcm.cu:15-33, adjlist.cu:62-87/:127-150, algo_bfs.cu:11-39.
#include <algorithm>
#include <numeric>

#include "graph.h"

namespace gcnb {
namespace graph {

// in-degree + out-degree of every vertex (Edgelist::compute_degrees, edgelist.cu:79-102)
static std::vector<int64_t> total_degree(const CsrView& g) {
  std::vector<int64_t> deg((size_t)g.m, 0);
  for (int64_t u = 0; u < g.m; ++u) deg[u] = g.rowptr[u + 1] - g.rowptr[u];
  for (int64_t z = 0; z < g.nnz; ++z) ++deg[g.col[z]];
  return deg;
}

std::vector<int64_t> order_deg(const CsrView& g, bool desc) {
  // Sort key: degree (descending or ascending), ties by vertex id ascending
  // (order_deg.cu:8-13).  Ids start in ascending order, so a stable counting sort on the
  // degree keeps the tie order; O(n + max_degree) instead of the reference's comparison sort.
  const std::vector<int64_t> deg = total_degree(g);
  const int64_t n = g.m;
  int64_t maxd = 0;
  for (int64_t d : deg) maxd = std::max(maxd, d);
  std::vector<int64_t> start((size_t)maxd + 2, 0);
  for (int64_t d : deg) ++start[(size_t)(desc ? maxd - d : d) + 1];
  for (size_t i = 1; i < start.size(); ++i) start[i] += start[i - 1];
  std::vector<int64_t> rank((size_t)n);
  for (int64_t u = 0; u < n; ++u) rank[u] = start[(size_t)(desc ? maxd - deg[u] : deg[u])]++;
  return rank;
}

FlatAdj relabelled_lists(const CsrView& g, const std::vector<int64_t>& rank, bool transpose) {
  const int64_t n = g.m;
  FlatAdj a;
  a.off.assign((size_t)n + 1, 0);
  a.adj.resize((size_t)g.nnz);
  for (int64_t u = 0; u < n; ++u)
    for (int64_t z = g.rowptr[u]; z < g.rowptr[u + 1]; ++z)
      ++a.off[(size_t)(transpose ? rank[g.col[z]] : rank[u]) + 1];
  for (int64_t u = 0; u < n; ++u) a.off[u + 1] += a.off[u];
  std::vector<int64_t> cur(a.off.begin(), a.off.end() - 1);
  for (int64_t u = 0; u < n; ++u) {
    const int64_t ru = rank[u];
    for (int64_t z = g.rowptr[u]; z < g.rowptr[u + 1]; ++z) {
      const int64_t rv = rank[g.col[z]];
      if (transpose) a.adj[(size_t)cur[rv]++] = (uint32_t)ru;
      else a.adj[(size_t)cur[ru]++] = (uint32_t)rv;
    }
  }
  parallel_rows(n, a.off, [&](int64_t u0, int64_t u1) {
    for (int64_t u = u0; u < u1; ++u) std::sort(a.adj.begin() + a.off[u], a.adj.begin() + a.off[u + 1]);
  });
  return a;
}

std::vector<uint32_t> bfs_order(const FlatAdj& out) {
  const int64_t n = (int64_t)out.off.size() - 1;
  std::vector<uint32_t> order;
  order.reserve((size_t)n);
  std::vector<uint8_t> placed((size_t)n, 0);
  size_t head = 0;
  for (int64_t root = 0; root < n; ++root) {
    if (placed[root]) continue;
    placed[root] = 1;
    order.push_back((uint32_t)root);
    while (head < order.size()) {
      const uint32_t w = order[head++];
      for (const uint32_t* p = out.begin(w); p != out.end(w); ++p) {
        if (placed[*p]) continue;
        placed[*p] = 1;
        order.push_back(*p);
      }
    }
  }
  return order;
}

std::vector<int64_t> order_rcm(const CsrView& g) {
  // order_rcm.cu:15-33 with directed = true: relabel by ascending degree, BFS over the
  // relabelled out-lists, reverse.
  const int64_t n = g.m;
  const std::vector<int64_t> rank_deg = order_deg(g, /*desc=*/false);
  const FlatAdj out = relabelled_lists(g, rank_deg, /*transpose=*/false);
  const std::vector<uint32_t> visit = bfs_order(out);
  std::vector<int64_t> pos((size_t)n);
  for (int64_t i = 0; i < n; ++i) pos[visit[i]] = i;
  std::vector<int64_t> rank((size_t)n);
  for (int64_t u = 0; u < n; ++u) rank[u] = n - 1 - pos[rank_deg[u]];
  return rank;
}

}  // namespace graph
}  // namespace gcnb
