// extern "C" entry points of the host graph preprocessing (include/gcn_b200.h) and the
// host-side part of the compatibility shim (dfs, gorder, perm_apply, csr2seg_Cmajor with the
// reference's names and argument orders, renumber.cu:23,157,233 and tile.cu:11-12).
#include <stdio.h>

#include <exception>
#include <new>
#include <vector>

#include "../common.h"
#include "graph.h"

using namespace gcnb;
using namespace gcnb::graph;

static int check_csr(const int32_t* rowptr, const int32_t* col, int64_t m, int64_t nnz) {
  GCNB_REQUIRE(m >= 0 && nnz >= 0, "negative size (m=%lld nnz=%lld)", (long long)m, (long long)nnz);
  GCNB_REQUIRE(rowptr != nullptr && (nnz == 0 || col != nullptr), "null CSR pointer");
  GCNB_REQUIRE(m < 0x7fffffffLL && nnz < 0x7fffffffLL, "int32 CSR limits exceeded");
  GCNB_REQUIRE(rowptr[0] == 0 && rowptr[m] == nnz, "rowptr[0]=%d rowptr[m]=%d, expected 0 and %lld",
               rowptr[0], rowptr[m], (long long)nnz);
  for (int64_t r = 0; r < m; ++r)
    GCNB_REQUIRE(rowptr[r + 1] >= rowptr[r], "rowptr decreases at row %lld", (long long)r);
  for (int64_t z = 0; z < nnz; ++z)
    GCNB_REQUIRE(col[z] >= 0 && col[z] < m, "column %d out of range at nz %lld (square graph expected)",
                 col[z], (long long)z);
  return GCNB_OK;
}

static int check_perm(const int32_t* vomp, int64_t m) {
  GCNB_REQUIRE(vomp != nullptr, "vomp is null");
  std::vector<uint8_t> seen((size_t)m, 0);
  for (int64_t i = 0; i < m; ++i) {
    GCNB_REQUIRE(vomp[i] >= 0 && vomp[i] < m && !seen[(size_t)vomp[i]],
                 "vomp is not a permutation (entry %lld = %d)", (long long)i, vomp[i]);
    seen[(size_t)vomp[i]] = 1;
  }
  return GCNB_OK;
}

// rank[old] = new  ->  vomp[new] = old  (renumber.cu:184-189)
static void invert_rank(const std::vector<int64_t>& rank, int32_t* vomp) {
  for (size_t old = 0; old < rank.size(); ++old) vomp[rank[old]] = (int32_t)old;
}

// C++ exceptions (std::bad_alloc from the work vectors, std::system_error from the host threads of
// parallel_rows) must not cross the C ABI: every entry point is a function-try-block.
#define GCNB_CATCH_ALL                                                                       \
  catch (const std::bad_alloc&) { return set_error(GCNB_ERR_ALLOC, "out of host memory"); }  \
  catch (const std::exception& e) { return set_error(GCNB_ERR_ALLOC, "host exception: %s", e.what()); } \
  catch (...) { return set_error(GCNB_ERR_ALLOC, "unknown host exception"); }

extern "C" {

int gcnb_order_deg(const int32_t* rowptr, const int32_t* col, int64_t m, int64_t nnz, int desc,
                   int64_t* rank_out) try {
  int rc = check_csr(rowptr, col, m, nnz);
  if (rc != GCNB_OK) return rc;
  GCNB_REQUIRE(rank_out != nullptr || m == 0, "rank_out is null");
  const std::vector<int64_t> r = order_deg({rowptr, col, m, nnz}, desc != 0);
  for (int64_t i = 0; i < m; ++i) rank_out[i] = r[i];
  return GCNB_OK;
} GCNB_CATCH_ALL

int gcnb_order_rcm(const int32_t* rowptr, const int32_t* col, int64_t m, int64_t nnz,
                   int64_t* rank_out) try {
  int rc = check_csr(rowptr, col, m, nnz);
  if (rc != GCNB_OK) return rc;
  GCNB_REQUIRE(rank_out != nullptr || m == 0, "rank_out is null");
  const std::vector<int64_t> r = order_rcm({rowptr, col, m, nnz});
  for (int64_t i = 0; i < m; ++i) rank_out[i] = r[i];
  return GCNB_OK;
} GCNB_CATCH_ALL

int gcnb_order_gorder(const int32_t* rowptr, const int32_t* col, int64_t m, int64_t nnz,
                      int window, int64_t* rank_out) try {
  int rc = check_csr(rowptr, col, m, nnz);
  if (rc != GCNB_OK) return rc;
  GCNB_REQUIRE(rank_out != nullptr || m == 0, "rank_out is null");
  GCNB_REQUIRE(window >= 1, "window = %d must be >= 1", window);
  std::vector<int64_t> r;
  if (!order_gorder({rowptr, col, m, nnz}, window, &r))
    return set_error(GCNB_ERR_UNSUPPORTED,
                     "Gorder needs every vertex to have an edge (add self-loops first); the "
                     "reference is undefined on isolated vertices (unitheap.cu:31-33)");
  for (int64_t i = 0; i < m; ++i) rank_out[i] = r[i];
  return GCNB_OK;
} GCNB_CATCH_ALL

int gcnb_perm_apply(int32_t* rowptr, int32_t* col, float* val, const int32_t* vomp, int64_t m,
                    int64_t nnz) try {
  int rc = check_csr(rowptr, col, m, nnz);
  if (rc != GCNB_OK) return rc;
  GCNB_REQUIRE(nnz == 0 || val != nullptr, "val is null");
  rc = check_perm(vomp, m);
  if (rc != GCNB_OK) return rc;
  perm_apply(rowptr, col, val, vomp, m, nnz);
  return GCNB_OK;
} GCNB_CATCH_ALL

static int renumber_with_rank(int32_t* rowptr, int32_t* col, float* val, int32_t* vomp, int64_t m,
                              int64_t nnz, const std::vector<int64_t>& rank) {
  GCNB_REQUIRE(vomp != nullptr || m == 0, "vomp is null");
  invert_rank(rank, vomp);
  perm_apply(rowptr, col, val, vomp, m, nnz);
  return GCNB_OK;
}

int gcnb_renumber_gorder(int32_t* rowptr, int32_t* col, float* val, int32_t* vomp, int64_t m,
                         int64_t nnz, int window) try {
  int rc = check_csr(rowptr, col, m, nnz);
  if (rc != GCNB_OK) return rc;
  GCNB_REQUIRE(window >= 1, "window = %d must be >= 1", window);
  std::vector<int64_t> r;
  if (!order_gorder({rowptr, col, m, nnz}, window, &r))
    return set_error(GCNB_ERR_UNSUPPORTED, "Gorder needs every vertex to have an edge");
  return renumber_with_rank(rowptr, col, val, vomp, m, nnz, r);
} GCNB_CATCH_ALL

int gcnb_renumber_rcm(int32_t* rowptr, int32_t* col, float* val, int32_t* vomp, int64_t m,
                      int64_t nnz) try {
  int rc = check_csr(rowptr, col, m, nnz);
  if (rc != GCNB_OK) return rc;
  return renumber_with_rank(rowptr, col, val, vomp, m, nnz, order_rcm({rowptr, col, m, nnz}));
} GCNB_CATCH_ALL

int gcnb_renumber_deg(int32_t* rowptr, int32_t* col, float* val, int32_t* vomp, int64_t m,
                      int64_t nnz, int desc) try {
  int rc = check_csr(rowptr, col, m, nnz);
  if (rc != GCNB_OK) return rc;
  return renumber_with_rank(rowptr, col, val, vomp, m, nnz,
                            order_deg({rowptr, col, m, nnz}, desc != 0));
} GCNB_CATCH_ALL

int gcnb_renumber_dfs(int32_t* rowptr, int32_t* col, float* val, int32_t* vomp, int64_t m,
                      int64_t nnz) try {
  int rc = check_csr(rowptr, col, m, nnz);
  if (rc != GCNB_OK) return rc;
  GCNB_REQUIRE(vomp != nullptr || m == 0, "vomp is null");
  dfs_order({rowptr, col, m, nnz}, vomp);
  perm_apply(rowptr, col, val, vomp, m, nnz);
  return GCNB_OK;
} GCNB_CATCH_ALL

int gcnb_renumber_rabbit(int32_t* rowptr, int32_t* col, float* val, int32_t* vomp, int64_t m,
                         int64_t nnz) try {
  int rc = check_csr(rowptr, col, m, nnz);
  if (rc != GCNB_OK) return rc;
  GCNB_REQUIRE(vomp != nullptr || m == 0, "vomp is null");
  rabbit_order({rowptr, col, m, nnz}, vomp);
  perm_apply(rowptr, col, val, vomp, m, nnz);
  return GCNB_OK;
} GCNB_CATCH_ALL

int gcnb_tile_count(const int32_t* rowptr, const int32_t* col, int64_t m, int64_t nnz, int tm,
                    int64_t* n_segs_out) try {
  GCNB_REQUIRE(rowptr && (nnz == 0 || col) && n_segs_out, "null pointer");
  GCNB_REQUIRE(tm >= 1 && tm <= 64, "tm = %d out of range", tm);
  GCNB_REQUIRE(m >= 0 && rowptr[m] == nnz, "rowptr[m] != nnz");
  *n_segs_out = build_segments(rowptr, col, nullptr, m, tm, 0, (m + tm - 1) / tm, 0, true, nullptr);
  return GCNB_OK;
} GCNB_CATCH_ALL

int gcnb_csr2tile(const int32_t* rowptr, const int32_t* col, const float* val, int64_t m,
                  int64_t n, int64_t nnz, int tm, int n_sm, int32_t* segVoMap,
                  int32_t* seg_rowPtr, float* segNzCV, int32_t* grouped_tailSeg,
                  int32_t* next_seg, int64_t* n_segs_out) try {
  (void)n;
  GCNB_REQUIRE(rowptr && (nnz == 0 || (col && val)) && n_segs_out, "null pointer");
  GCNB_REQUIRE(segVoMap && seg_rowPtr && (nnz == 0 || segNzCV), "null output pointer");
  GCNB_REQUIRE(tm >= 1 && tm <= 64, "tm = %d out of range", tm);
  GCNB_REQUIRE(m >= 0 && rowptr[m] == nnz, "rowptr[m] != nnz");
  GCNB_REQUIRE((grouped_tailSeg == nullptr) == (next_seg == nullptr),
               "grouped_tailSeg and next_seg go together");
  GCNB_REQUIRE(next_seg == nullptr || n_sm >= 1, "n_sm = %d must be >= 1", n_sm);
  TileOut out{segVoMap, seg_rowPtr, segNzCV, grouped_tailSeg, next_seg};
  const int64_t n_segs =
      build_segments(rowptr, col, val, m, tm, 0, (m + tm - 1) / tm, 0, true, &out);
  if (next_seg) build_buckets(seg_rowPtr, n_segs, nnz, tm, n_sm, grouped_tailSeg, next_seg);
  *n_segs_out = n_segs;
  return GCNB_OK;
} GCNB_CATCH_ALL

// ---- compat: reference names (void, errors to stderr + gcnb_last_error) -------------------------
static void report(const char* fn, int rc) {
  if (rc != GCNB_OK) fprintf(stderr, "gcn_b200: %s failed (%d): %s\n", fn, rc, gcnb_last_error());
}

void dfs(int* rowPtr, int* col, float* vals, int* vomp, int m, int n, int nnz) {
  (void)n;
  report("dfs", gcnb_renumber_dfs(rowPtr, col, vals, vomp, m, nnz));
}

void rabbit(int* rowPtr, int* col, float* vals, int* vomp, int m, int n, int nnz) {
  (void)n;
  report("rabbit", gcnb_renumber_rabbit(rowPtr, col, vals, vomp, m, nnz));
}

void gorder(int* rowPtr, int* col, float* vals, int* vomp, int m, int n, int nnz) {
  (void)n;
  report("gorder", gcnb_renumber_gorder(rowPtr, col, vals, vomp, m, nnz, /*window=*/3));  // renumber.cu:176
}

void perm_apply(int* rowPtr, int* col, float* vals, int* vomp, int m, int n, int nnz) {
  (void)n;
  report("perm_apply", gcnb_perm_apply(rowPtr, col, vals, vomp, m, nnz));
}

void csr2seg_Cmajor(int ridx, int* rowPtr, int* colIdx, float* vals, int m, int n, int nnz,
                    int* voMp, int* segVoMap, int* seg_rowPtr, float* segNzCV, int tm,
                    int* n_segs) {
  (void)n; (void)nnz; (void)voMp;
  if (!rowPtr || !colIdx || !vals || !segVoMap || !seg_rowPtr || !segNzCV || !n_segs || tm < 1) {
    set_error(GCNB_ERR_ARG, "csr2seg_Cmajor: null pointer or bad tm");
    report("csr2seg_Cmajor", GCNB_ERR_ARG);
    return;
  }
  TileOut out{segVoMap, seg_rowPtr, segNzCV, nullptr, nullptr};
  n_segs[0] = (int)build_segments(rowPtr, colIdx, vals, m, tm, ridx, ridx + 1, n_segs[0], true, &out);
}

}  // extern "C"
