// Rabbit order: modularity-driven agglomeration + dendrogram leaf order (the reference's default
// reordering, RBT = 1 in pygcn/gcn6.py:27-30; algorithm of renumber.cu:319-522 with
// opt_iterative = true, hub grouping off, shyness 1).
//
// Written from the behaviour of the reference so that vomp is bit-exact with it, but on flat data:
// the per-cluster neighbour weights are open-addressing hash tables instead of one std::map per
// vertex (merging into a hub is O(1) per edge; on the Reddit-shaped graph sorted vectors were
// quadratic), the dendrogram is two index arrays instead of pointer nodes (and is walked iteratively —
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

// Neighbour weights of one cluster: open-addressing hash table (linear probing, backward-shift
// deletion, power-of-two capacity, load <= 1/2).  The reference keeps a std::map per vertex and so
// visits neighbours in ascending id; the only place where the order is observable is the choice of
// the merge target — "first strict maximum in ascending id" — which is the same as "maximum, ties to
// the smallest id" and needs no order.  Everything else adds integers, which commutes.
struct Table {
  std::vector<int32_t> key, val;   // key < 0: empty slot
  int32_t count = 0;
  uint32_t mask = 0;

  static uint32_t hash(int32_t k) { return (uint32_t)k * 2654435761u; }
  void init(int32_t expected) {
    uint32_t cap = 4;
    while (cap < 2u * (uint32_t)expected + 2u) cap <<= 1;
    key.assign(cap, -1);
    val.assign(cap, 0);
    mask = cap - 1;
    count = 0;
  }
  void grow() {
    std::vector<int32_t> ok, ov;
    ok.swap(key);
    ov.swap(val);
    const uint32_t cap = (uint32_t)ok.size() * 2;
    key.assign(cap, -1);
    val.assign(cap, 0);
    mask = cap - 1;
    count = 0;
    for (size_t i = 0; i < ok.size(); ++i)
      if (ok[i] >= 0) add(ok[i], ov[i]);
  }
  int32_t* find(int32_t k) {
    for (uint32_t i = hash(k) & mask;; i = (i + 1) & mask) {
      if (key[i] == k) return &val[i];
      if (key[i] < 0) return nullptr;
    }
  }
  void add(int32_t k, int32_t w) {   // table[k] += w
    if (2u * (uint32_t)(count + 1) > mask + 1) grow();
    for (uint32_t i = hash(k) & mask;; i = (i + 1) & mask) {
      if (key[i] == k) { val[i] += w; return; }
      if (key[i] < 0) { key[i] = k; val[i] = w; ++count; return; }
    }
  }
  bool erase(int32_t k, int32_t* w_out) {
    uint32_t i = hash(k) & mask;
    for (;; i = (i + 1) & mask) {
      if (key[i] == k) break;
      if (key[i] < 0) return false;
    }
    *w_out = val[i];
    // backward-shift deletion keeps every probe sequence gap-free
    for (uint32_t j = (i + 1) & mask;; j = (j + 1) & mask) {
      if (key[j] < 0) break;
      const uint32_t home = hash(key[j]) & mask;
      if (((j - home) & mask) >= ((j - i) & mask)) {
        key[i] = key[j];
        val[i] = val[j];
        i = j;
      }
    }
    key[i] = -1;
    --count;
    return true;
  }
  void release() {
    std::vector<int32_t>().swap(key);
    std::vector<int32_t>().swap(val);
    count = 0;
    mask = 0;
  }
};

}  // namespace

void rabbit_order(const CsrView& g, int32_t* vomp) {
  const int64_t n = g.m;
  if (n == 0) return;
  // undirected unit-weight neighbour tables without self loops
  std::vector<Table> nbr((size_t)n);
  std::vector<int32_t> deg((size_t)n, 0);
  {
    std::vector<int32_t> cnt((size_t)n, 0);
    for (int64_t v = 0; v < n; ++v)
      for (int64_t z = g.rowptr[v]; z < g.rowptr[v + 1]; ++z) {
        const int32_t d = g.col[z];
        if (d != v) { ++cnt[(size_t)v]; ++cnt[(size_t)d]; }
      }
    for (int64_t v = 0; v < n; ++v) nbr[(size_t)v].init(cnt[(size_t)v]);
    for (int64_t v = 0; v < n; ++v)
      for (int64_t z = g.rowptr[v]; z < g.rowptr[v + 1]; ++z) {
        const int32_t d = g.col[z];
        if (d == v) continue;
        if (!nbr[(size_t)v].find(d)) nbr[(size_t)v].add(d, 1);
        if (!nbr[(size_t)d].find((int32_t)v)) nbr[(size_t)d].add((int32_t)v, 1);
      }
    // degree as the reference sees it while it initialises vertex v: out-neighbours of v plus the
    // in-neighbours that were initialised before it
    std::vector<uint8_t> is_out((size_t)n, 0);
    for (int64_t v = 0; v < n; ++v) {
      for (int64_t z = g.rowptr[v]; z < g.rowptr[v + 1]; ++z) is_out[(size_t)g.col[z]] = 1;
      int32_t dv = 0;
      const Table& t = nbr[(size_t)v];
      for (size_t i = 0; i < t.key.size(); ++i)
        if (t.key[i] >= 0 && (is_out[(size_t)t.key[i]] || t.key[i] < v)) ++dv;
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
      Table& ru = nbr[u];
      double best = -1;
      int32_t v = -1;
      const double dv_2m = deg[u] * two_m_inv;
      for (size_t i = 0; i < ru.key.size(); ++i) {
        const int32_t d = ru.key[i];
        if (d < 0) continue;
        const double dq = ru.val[i] - deg[(size_t)d] * dv_2m;
        // first strict maximum in ascending id == maximum with ties to the smallest id
        if (dq > best || (dq == best && v >= 0 && d < v)) { best = dq; v = d; }
      }
      if (best <= 0) continue;

      // u joins v
      deg[(size_t)v] += deg[u];
      Table& rv = nbr[(size_t)v];
      for (size_t i = 0; i < ru.key.size(); ++i) {
        const int32_t d = ru.key[i];
        if (d < 0 || d == v) continue;
        rv.add(d, ru.val[i]);
        Table& rd = nbr[(size_t)d];
        int32_t w = 0;
        if (!rd.erase((int32_t)u, &w)) continue;
        rd.add(v, w);
      }
      {
        int32_t w = 0;
        rv.erase((int32_t)u, &w);
      }
      left[u] = tree[(size_t)v];
      right[u] = tree[u];
      tree[u] = -1;
      tree[(size_t)v] = (int32_t)(n + u);
      ru.release();   // never read again

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
