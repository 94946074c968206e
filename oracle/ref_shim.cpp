// TEST INFRASTRUCTURE ONLY — not part of the shipped product.
//
// Thin C-ABI shim over the *unmodified* reference sources so that the reference's
// C++-linkage ordering functions can be called from the parity tests.  It is compiled
// together with the reference's own files where they lie under /root/reference
// (see oracle/Makefile, target `ref`); nothing from the reference is copied here.
//
//   ref_order_deg    -> order_deg(Edgelist&, bool desc)        order_deg.cu:41-45
//   ref_order_rcm    -> order_rcm(Edgelist&, bool directed)    order_rcm.cu:15-33
//   ref_order_gorder -> complete_gorder(Edgelist&, ul window)  order_gorder.cu:13-31
//
// All three return rank[old] = new (the reference's convention, order_rcm.cu:28-31);
// gorder() in renumber.cu:184-189 inverts that into vomp[new] = old.
#include <vector>
#include "tools.cuh"
#include "edgelist.cuh"
#include "order_deg.cuh"
#include "order_rcm.cuh"
#include "order_gorder.cuh"

extern "C" {

void ref_order_deg(int* rowPtr, int* col, int m, int nnz, int desc, long long* rank_out) {
  Edgelist h(rowPtr, col, m, nnz);
  std::vector<ul> r = order_deg(h, desc != 0);
  for (int i = 0; i < m; ++i) rank_out[i] = (long long)r[i];
}

void ref_order_rcm(int* rowPtr, int* col, int m, int nnz, long long* rank_out) {
  Edgelist h(rowPtr, col, m, nnz);
  std::vector<ul> r = order_rcm(h, true);
  for (int i = 0; i < m; ++i) rank_out[i] = (long long)r[i];
}

void ref_order_gorder(int* rowPtr, int* col, int m, int nnz, int window, long long* rank_out) {
  Edgelist h(rowPtr, col, m, nnz);
  std::vector<unsigned int> r = complete_gorder(h, (ul)window);
  for (int i = 0; i < m; ++i) rank_out[i] = (long long)r[i];
}

}  // extern "C"
