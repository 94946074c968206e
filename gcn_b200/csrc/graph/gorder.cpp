// Gorder (Wei, Yu, Lu, Lin, SIGMOD'16) with the exact pick order of the reference's
// order_gorder.cu:13-143 + unitheap.cu:16-217: a list of vertices kept in score-descending
// order with one (head, tail) pair per score value and lazily applied negative updates.
// Own implementation on flat arrays; the observable order (which vertex ExtractMax returns at
// every step, including all tie-breaks) is what is reproduced.
#include <algorithm>
#include <climits>
#include <cmath>

#include "graph.h"

namespace gcnb {
namespace graph {

namespace {

class ScoreList {
 public:
  static constexpr int kFrozen = INT_MAX / 2;  // "not in the heap" marker (unitheap.cuh:37)

  explicit ScoreList(int64_t n)
      : nil_((uint32_t)n + 2), key_((size_t)n, kFrozen), pending_((size_t)n, kFrozen),
        prev_((size_t)n, nil_), next_((size_t)n, nil_), top_(nil_) {}

  uint32_t nil() const { return nil_; }
  uint32_t top() const { return top_; }
  int64_t size() const { return size_; }

  // unitheap.cu:23-28: the key is entered with an equal pending decrement, i.e. the
  // effective score starts at 0 and rises through lazy_add(+1)
  void insert(uint32_t v, int key) {
    key_[v] = key;
    pending_[v] = -key;
    ++size_;
  }

  // unitheap.cu:30-62: link vertices 0..size-1 by key descending, id ascending
  void build() {
    const int64_t n = size_;
    int maxk = 0;
    for (int64_t v = 0; v < n; ++v) maxk = std::max(maxk, key_[v]);
    // stable counting sort by descending key
    std::vector<int64_t> start((size_t)maxk + 2, 0);
    for (int64_t v = 0; v < n; ++v) ++start[(size_t)(maxk - key_[v]) + 1];
    for (size_t i = 1; i < start.size(); ++i) start[i] += start[i - 1];
    std::vector<uint32_t> seq((size_t)n);
    for (int64_t v = 0; v < n; ++v) seq[(size_t)start[(size_t)(maxk - key_[v])]++] = (uint32_t)v;
    head_.assign((size_t)maxk + 1, nil_);
    tail_.assign((size_t)maxk + 1, nil_);
    for (int64_t i = 0; i < n; ++i) {
      const uint32_t v = seq[i];
      prev_[v] = i > 0 ? seq[i - 1] : nil_;
      next_[v] = i + 1 < n ? seq[i + 1] : nil_;
      const int k = key_[v];
      if (head_[k] == nil_) head_[k] = v;
      tail_[k] = v;
    }
    top_ = n > 0 ? seq[0] : nil_;
  }

  // unitheap.cu:152-170
  void remove(uint32_t v) {
    pending_[v] = kFrozen;
    const uint32_t p = prev_[v], x = next_[v];
    if (p != nil_) next_[p] = x;
    if (x != nil_) prev_[x] = p;
    unlink_from_level(v, x, p);
    if (top_ == v) top_ = x;
    prev_[v] = next_[v] = nil_;
    --size_;
  }

  // unitheap.cu:82-96: settle the pending decrement of the head until the head is stable
  uint32_t pop_max() {
    uint32_t seen;
    do {
      seen = top_;
      if (pending_[top_] < 0) sink_top();
    } while (top_ != seen);
    remove(top_);
    return seen;
  }

  // unitheap.cu:177-185
  void lazy_add(uint32_t v, int up) {
    if (pending_[v] == kFrozen) return;
    if (pending_[v] == 0 && up > 0) raise(v);
    else pending_[v] += up;
  }

 private:
  void grow_levels(int k) {
    if ((size_t)k >= head_.size()) {
      head_.resize((size_t)k + 1, nil_);
      tail_.resize((size_t)k + 1, nil_);
    }
  }

  // unitheap.cu:68-76
  void unlink_from_level(uint32_t v, uint32_t nxt, uint32_t prv) {
    const int k = key_[v];
    if (head_[k] == tail_[k]) head_[k] = tail_[k] = nil_;
    else if (v == head_[k]) head_[k] = nxt;
    else if (v == tail_[k]) tail_[k] = prv;
  }

  // unitheap.cu:99-149: apply half (rounded away from the list) of the head's pending
  // decrement and move it behind every vertex whose key is >= its new key
  void sink_top() {
    const uint32_t t = top_;
    const uint32_t nxt = next_[t];
    if (nxt == nil_) return;
    const int key = key_[t];
    const int leftover = pending_[t] / 2;  // C division: truncates toward zero
    const int new_key = key + pending_[t] - leftover;
    if (new_key >= key_[nxt]) return;
    pending_[t] = leftover;
    uint32_t level_tail = tail_[key];
    uint32_t below = next_[level_tail];
    while (below != nil_ && key_[below] >= new_key) {
      level_tail = tail_[key_[below]];
      below = next_[level_tail];
    }
    prev_[nxt] = nil_;
    prev_[t] = level_tail;
    next_[t] = below;
    next_[level_tail] = t;
    if (below != nil_) prev_[below] = t;
    unlink_from_level(t, nxt, nil_);
    key_[t] = new_key;
    tail_[new_key] = t;
    if (head_[new_key] == nil_) head_[new_key] = t;
    top_ = nxt;
  }

  // unitheap.cu:187-217: key += 1, moving the vertex to the front of its old level
  void raise(uint32_t v) {
    const uint32_t level_head = head_[key_[v]];
    const uint32_t p = prev_[v], x = next_[v];
    if (level_head != v) {
      next_[p] = x;
      if (x != nil_) prev_[x] = p;
      const uint32_t above = prev_[level_head];
      prev_[v] = above;
      next_[v] = level_head;
      prev_[level_head] = v;
      if (above != nil_) next_[above] = v;
    }
    unlink_from_level(v, x, p);
    const int k = ++key_[v];
    grow_levels(k);
    tail_[k] = v;
    if (head_[k] == nil_) {
      head_[k] = v;
      if (k > key_[top_]) top_ = v;
    }
  }

  uint32_t nil_;
  std::vector<int> key_, pending_;
  std::vector<uint32_t> prev_, next_;
  std::vector<uint32_t> head_, tail_;
  uint32_t top_;
  int64_t size_ = 0;
};

// order_gorder.cu:88-143.  `entering` joins the window, `leaving` drops out of it (equal on
// the first calls, meaning nothing leaves).  Scores: +1/-1 to children, to parents and to the
// parents' other children, skipping hubs with more than `hub` out-neighbours; parents common
// to both vertices cancel and are skipped (sorted-list merge).
void slide_window(const FlatAdj& out, const FlatAdj& in, ScoreList& heap, int64_t hub,
                  uint32_t entering, uint32_t leaving, std::vector<uint32_t>& scratch_old,
                  std::vector<uint32_t>& scratch_new) {
  const uint32_t* po = in.begin(leaving);
  const uint32_t* po_end = in.end(leaving);
  const uint32_t* pn = in.begin(entering);
  const uint32_t* pn_end = in.end(entering);
  if (leaving == entering) {
    po = po_end;
  } else if (out.degree(leaving) <= hub) {
    for (const uint32_t* c = out.begin(leaving); c != out.end(leaving); ++c) heap.lazy_add(*c, -1);
  }
  scratch_old.clear();
  scratch_new.clear();
  while (po != po_end || pn != pn_end) {
    bool take_new;
    if (po == po_end) take_new = true;
    else if (pn == pn_end) take_new = false;
    else if (*pn == *po) { ++pn; ++po; continue; }
    else take_new = *pn < *po;
    if (take_new) {
      if (out.degree(*pn) <= hub) scratch_new.push_back(*pn);
      ++pn;
    } else {
      if (out.degree(*po) <= hub) scratch_old.push_back(*po);
      ++po;
    }
  }
  for (uint32_t parent : scratch_old) {
    heap.lazy_add(parent, -1);
    for (const uint32_t* s = out.begin(parent); s != out.end(parent); ++s)
      if (*s != leaving) heap.lazy_add(*s, -1);
  }
  if (out.degree(entering) <= hub)
    for (const uint32_t* c = out.begin(entering); c != out.end(entering); ++c) heap.lazy_add(*c, +1);
  for (uint32_t parent : scratch_new) {
    heap.lazy_add(parent, +1);
    for (const uint32_t* s = out.begin(parent); s != out.end(parent); ++s)
      if (*s != entering) heap.lazy_add(*s, +1);
  }
}

}  // namespace

std::vector<int64_t> gorder_core(const FlatAdj& out, const FlatAdj& in, int window) {
  const int64_t n = (int64_t)out.off.size() - 1;
  std::vector<int64_t> rank((size_t)n);
  if (n == 0) return rank;
  ScoreList heap(n);
  const int64_t hub = (int64_t)std::sqrt((double)n);  // unitheap.cu:19
  for (int64_t u = 0; u < n; ++u) heap.insert((uint32_t)u, (int)in.degree(u));
  heap.build();
  std::vector<uint32_t> order;
  order.reserve((size_t)n);
  std::vector<uint32_t> so, sn;
  const uint32_t first = heap.top();
  order.push_back(first);
  heap.remove(first);
  slide_window(out, in, heap, hub, first, first, so, sn);
  while (heap.size() > 0) {
    const uint32_t v = heap.pop_max();
    order.push_back(v);
    uint32_t leaving = v;
    if ((int64_t)order.size() > window) leaving = order[order.size() - (size_t)window - 1];
    slide_window(out, in, heap, hub, v, leaving, so, sn);
  }
  for (int64_t i = 0; i < n; ++i) rank[order[i]] = i;
  return rank;
}

bool order_gorder(const CsrView& g, int window, std::vector<int64_t>* rank) {
  // complete_gorder, order_gorder.cu:13-31: RCM first, Gorder on the RCM-relabelled graph,
  // ranks composed.
  const int64_t n = g.m;
  rank->clear();
  const std::vector<int64_t> rcm = order_rcm(g);
  const FlatAdj out = relabelled_lists(g, rcm, false);
  const FlatAdj in = relabelled_lists(g, rcm, true);
  for (int64_t u = 0; u < n; ++u)
    if (out.degree(u) + in.degree(u) == 0) return false;
  const std::vector<int64_t> gr = gorder_core(out, in, window);
  rank->resize((size_t)n);
  for (int64_t u = 0; u < n; ++u) (*rank)[u] = gr[rcm[u]];
  return true;
}

}  // namespace graph
}  // namespace gcnb
