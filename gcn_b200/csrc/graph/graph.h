// Host-side graph preprocessing of libgcn_b200.so: vertex orders (degree, RCM, Gorder), CSR
// renumbering and the reference tile format.  Written from the behaviour of the reference
// (order_deg.cu, order_rcm.cu, order_gorder.cu, unitheap.cu, adjlist.cu, algo_bfs.cu,
// renumber.cu, tile.cu) so that every permutation and every tile array is bit-exact with it,
// but on flat CSR arrays: no edge-list of pairs, no per-vertex std::vector, no std::map.
#pragma once
#include <stdint.h>
#include <stdlib.h>

#include <algorithm>
#include <thread>
#include <vector>

namespace gcnb {
namespace graph {

// Host threads for the loops whose iterations are independent (one row each): the hardware's, at most
// 16, or GCNB_HOST_THREADS.  The per-row work is unchanged, so the results are.
inline int host_threads() {
  const char* e = getenv("GCNB_HOST_THREADS");
  int t = e ? atoi(e) : (int)std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
  return std::max(1, t);
}

// fn(begin, end) over [0, n) cut into contiguous chunks of about equal weight, where
// prefix[i+1] - prefix[i] is the weight of item i (a CSR row pointer); one thread per chunk.
template <typename Prefix, typename Fn>
void parallel_rows(int64_t n, const Prefix& prefix, Fn fn) {
  const int64_t total = (int64_t)prefix[n] - (int64_t)prefix[0];
  const int t = (int)std::min<int64_t>(host_threads(), std::max<int64_t>(1, total / (1 << 16)));
  if (t <= 1 || n < 2) {
    fn((int64_t)0, n);
    return;
  }
  std::vector<int64_t> cut((size_t)t + 1, n);
  cut[0] = 0;
  int64_t i = 0;
  for (int c = 1; c < t; ++c) {
    const int64_t want = (int64_t)prefix[0] + total * c / t;
    while (i < n && (int64_t)prefix[i] < want) ++i;
    cut[(size_t)c] = i;
  }
  std::vector<std::thread> pool;
  for (int c = 1; c < t; ++c)
    if (cut[(size_t)c] < cut[(size_t)c + 1]) pool.emplace_back(fn, cut[(size_t)c], cut[(size_t)c + 1]);
  fn(cut[0], cut[1]);
  for (auto& th : pool) th.join();
}

struct CsrView {
  const int32_t* rowptr;
  const int32_t* col;
  int64_t m;
  int64_t nnz;
};

// Flat adjacency over relabelled vertex ids: neighbours of u are adj[off[u] .. off[u+1]),
// ascending.  (What the reference's Dadjlist/Badjlist hold after build_from_edgelist_ranked +
// sort_neighbours, adjlist.cu:62-87, :127-187.)
struct FlatAdj {
  std::vector<int64_t> off;
  std::vector<uint32_t> adj;
  int64_t degree(int64_t u) const { return off[u + 1] - off[u]; }
  const uint32_t* begin(int64_t u) const { return adj.data() + off[u]; }
  const uint32_t* end(int64_t u) const { return adj.data() + off[u + 1]; }
};

// rank[old] = new for all three (order_rcm.cu:28-31)
std::vector<int64_t> order_deg(const CsrView& g, bool desc);
std::vector<int64_t> order_rcm(const CsrView& g);
// returns false if the graph has an isolated vertex (the reference's heap then reads
// uninitialised links, unitheap.cu:31-33); rank is left empty
bool order_gorder(const CsrView& g, int window, std::vector<int64_t>* rank);

// out-lists (transpose = false) or in-lists (true) of the graph relabelled by rank
FlatAdj relabelled_lists(const CsrView& g, const std::vector<int64_t>& rank, bool transpose);
// BFS visit order over out-lists, restarting at the lowest unplaced id (algo_bfs.cu:11-39)
std::vector<uint32_t> bfs_order(const FlatAdj& out);
// Gorder proper on a relabelled graph (order_gorder.cu:35-84)
std::vector<int64_t> gorder_core(const FlatAdj& out, const FlatAdj& in, int window);

// P·A·P^T in place, vomp[new] = old (renumber.cu:233-318)
void perm_apply(int32_t* rowptr, int32_t* col, float* val, const int32_t* vomp, int64_t m,
                int64_t nnz);
// DFS pre-order from vertex 0, restart at lowest unvisited (renumber.cu:56-95): vomp[new]=old
void dfs_order(const CsrView& g, int32_t* vomp);

// Rabbit order (renumber.cu:319-522: iterative modularity agglomeration, dendrogram leaf order,
// hub grouping off): vomp[new] = old
void rabbit_order(const CsrView& g, int32_t* vomp);

// Reference tile format (tile.cu:11-169)
struct TileOut {
  int32_t* segVoMap;         // [tm * n_segs]        may be null (count only)
  int32_t* seg_rowPtr;       // [(tm+1) * n_segs]
  float* segNzCV;            // [2 * nnz]
  int32_t* grouped_tailSeg;  // [n_sm + 1]           may be null
  int32_t* next_seg;         // [n_sm + 1]           may be null
};
// Walks every tm-row panel in column order and closes a segment at >= 116 non-zeros or at the
// panel's last occupied column.  start_seg/first_panel/last_panel let the compat
// csr2seg_Cmajor process one panel; cumulative selects the corrected seg_rowPtr.
int64_t build_segments(const int32_t* rowptr, const int32_t* col, const float* val, int64_t m,
                       int tm, int64_t first_panel, int64_t last_panel, int64_t start_seg,
                       bool cumulative, const TileOut* out);
void build_buckets(const int32_t* seg_rowPtr, int64_t n_segs, int64_t nnz, int tm, int n_sm,
                   int32_t* grouped_tailSeg, int32_t* next_seg);

}  // namespace graph
}  // namespace gcnb
