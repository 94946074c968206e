// Rabbit order: modularity-driven agglomeration + dendrogram leaf order (the reference's default
// reordering, RBT = 1 in pygcn/gcn6.py:27-30; algorithm of renumber.cu:319-522 with
// opt_iterative = true, hub grouping off, shyness 1).
//
// Written from the behaviour of the reference so that vomp is bit-exact with it, but on flat data:
// the per-cluster neighbour weights are sorted (key, weight) vectors instead of one std::map per
// vertex, the dendrogram is two index arrays instead of pointer nodes (and is walked iteratively —
// the reference recurses, which overflows the stack on path-like merges), degrees and rounds are
// plain arrays.  What has to match exactly for the permutation to match:
//   * a cluster's neighbours are visited in ascending id and the FIRST strict maximum of
//     dQ = w - deg[d] * (deg[u] / 2m) wins (set_max, renumber.cu:15-20), evaluated in double in that
//     association;
//   * every round visits the surviving vertices in the order std::sort (introsort, not stable) leaves
//     them when keyed by current cluster degree (renumber.cu:407-408) — same library routine, same
//     input order, same result;
//   * deg[v] is taken when v is initialised, i.e. it counts v's out-neighbours and only those
//     in-neighbours with a smaller id (renumber.cu:383-399) — identical to the full degree on the
//     symmetric matrices GCN uses, reproduced anyway.
#include <algorithm>
#include <utility>

#include "graph.h"

namespace gcnb {
namespace graph {

namespace {

using Entry = std::pair<int32_t, int32_t>;   // (neighbour cluster, accumulated weight)
using Row = std::vector<Entry>;

inline Row::iterator find_key(Row& r, int32_t key) {
  return std::lower_bound(r.begin(), r.end(), key,
                          [](const Entry& e, int32_t k) { return e.first < k; });
}

// r[key] += w (inserting the key when absent), keeping r sorted
inline void add_weight(Row& r, int32_t key, int32_t w) {
  auto it = find_key(r, key);
  if (it != r.end() && it->first == key) it->second += w;
  else r.insert(it, Entry(key, w));
}

}  // namespace

void rabbit_order(const CsrView& g, int32_t* vomp) {
  const int64_t n = g.m;
  if (n == 0) return;
  // undirected unit-weight neighbour lists without self loops
  std::vector<Row> nbr((size_t)n);
  std::vector<int32_t> deg((size_t)n, 0);
  {
    std::vector<int32_t> cnt((size_t)n, 0);
    for (int64_t v = 0; v < n; ++v)
      for (int64_t z = g.rowptr[v]; z < g.rowptr[v + 1]; ++z) {
        const int32_t d = g.col[z];
        if (d != v) { ++cnt[(size_t)v]; ++cnt[(size_t)d]; }
      }
    for (int64_t v = 0; v < n; ++v) nbr[(size_t)v].reserve((size_t)cnt[(size_t)v]);
    for (int64_t v = 0; v < n; ++v)
      for (int64_t z = g.rowptr[v]; z < g.rowptr[v + 1]; ++z) {
        const int32_t d = g.col[z];
        if (d == v) continue;
        nbr[(size_t)v].emplace_back(d, 1);
        nbr[(size_t)d].emplace_back((int32_t)v, 1);
      }
    for (int64_t v = 0; v < n; ++v) {
      Row& r = nbr[(size_t)v];
      std::sort(r.begin(), r.end());
      r.erase(std::unique(r.begin(), r.end(), [](const Entry& a, const Entry& b) { return a.first == b.first; }),
              r.end());
    }
    // degree as the reference sees it while it initialises vertex v: out-neighbours of v plus the
    // in-neighbours that were initialised before it
    std::vector<uint8_t> is_out((size_t)n, 0);
    for (int64_t v = 0; v < n; ++v) {
      for (int64_t z = g.rowptr[v]; z < g.rowptr[v + 1]; ++z) is_out[(size_t)g.col[z]] = 1;
      int32_t dv = 0;
      for (const Entry& e : nbr[(size_t)v])
        if (is_out[(size_t)e.first] || e.first < v) ++dv;
      deg[(size_t)v] = dv;
      for (int64_t z = g.rowptr[v]; z < g.rowptr[v + 1]; ++z) is_out[(size_t)g.col[z]] = 0;
    }
  }
  int n_edges = 0;   // `int` like the reference: 2 * n_edges below is an int product
  for (int64_t v = 0; v < n; ++v) n_edges += deg[(size_t)v];
  const double two_m_inv = 1.0 / double(2 * n_edges);

  // dendrogram: node ids 0..n-1 are leaves, n+u is the cluster node created when u is merged away
  std::vector<int32_t> left((size_t)n, -1), right((size_t)n, -1);   // indexed by u
  std::vector<int32_t> tree((size_t)n);                            // root node of the cluster led by v, -1 if merged
  std::vector<int32_t> round_of((size_t)n, 0);
  std::vector<uint32_t> cur((size_t)n), next;
  for (int64_t v = 0; v < n; ++v) {
    tree[(size_t)v] = (int32_t)v;
    cur[(size_t)v] = (uint32_t)v;
  }

  for (int round = 1; !cur.empty(); ++round) {
    std::sort(cur.begin(), cur.end(),
              [&](uint32_t a, uint32_t b) { return deg[a] < deg[b]; });
    for (const uint32_t u : cur) {
      if (round_of[u] == round) continue;   // u absorbed somebody this round: it waits for the next
      Row& ru = nbr[u];
      double best = -1;
      int32_t v = -1;
      const double dv_2m = deg[u] * two_m_inv;
      for (const Entry& e : ru) {
        const double dq = e.second - deg[(size_t)e.first] * dv_2m;
        if (dq > best) { best = dq; v = e.first; }
      }
      if (best <= 0) continue;

      // u joins v
      deg[(size_t)v] += deg[u];
      Row& rv = nbr[(size_t)v];
      for (const Entry& e : ru) {
        const int32_t d = e.first;
        if (d == v) continue;
        add_weight(rv, d, e.second);
        Row& rd = nbr[(size_t)d];
        auto it = find_key(rd, (int32_t)u);
        if (it == rd.end() || it->first != (int32_t)u) continue;
        const int32_t w = it->second;
        rd.erase(it);
        add_weight(rd, v, w);
      }
      {
        auto it = find_key(rv, (int32_t)u);
        if (it != rv.end() && it->first == (int32_t)u) rv.erase(it);
      }
      left[u] = tree[(size_t)v];
      right[u] = tree[u];
      tree[u] = -1;
      tree[(size_t)v] = (int32_t)(n + u);
      Row().swap(ru);   // never read again

      if (round_of[(size_t)v] == round) continue;
      round_of[(size_t)v] = round;
      next.push_back((uint32_t)v);
    }
    cur.swap(next);
    next.clear();
  }

  // leaves of every surviving cluster, clusters in ascending leader id, left subtree first
  int64_t placed = 0;
  std::vector<int32_t> stack;
  for (int64_t v = 0; v < n; ++v) {
    if (tree[(size_t)v] < 0) continue;
    stack.push_back(tree[(size_t)v]);
    while (!stack.empty()) {
      const int32_t node = stack.back();
      stack.pop_back();
      if (node < n) {
        vomp[placed++] = node;
      } else {
        stack.push_back(right[(size_t)(node - n)]);
        stack.push_back(left[(size_t)(node - n)]);
      }
    }
  }
}

}  // namespace graph
}  // namespace gcnb
