// Compatibility shim, device side: the reference's exported names with their argument orders
// and `void` returns (flexspmm.cu:499-502, cuspmm.cu:23-24, tile.cu:104-106, permutate.cu:40-41),
// so the reference's ctypes call sites (pygcn/gcn6.py:39,54,334-352,374) can bind this library
// unchanged.  Launches go to the legacy default stream like the reference's `<<<g,b>>>`.
#include <stdio.h>

#include "common.h"
#include "graph/graph.h"

namespace gcnb {

// One warp per tile-segment of the reference format: tm rows, each row's pairs contiguous in
// segNzCV with the column id stored as a float (tile.cu:62-73).  Rows whose MSB is set in
// segVoMap are shared with other segments and are accumulated (the format's contract,
// flexspmm.cu:65,85-89 — the only place in this library where an atomic is needed, because the
// reference format splits rows across independent segments); other rows are stored.
// Lanes own interleaved columns, 4 per pass, so B rows are read with coalesced 128-byte
// requests; there is no per-SM bucket walk and no mutable next_seg state.
__global__ void __launch_bounds__(128)
flex_tiles_kernel(const int32_t* __restrict__ seg_rowPtr, const float* __restrict__ segNzCV,
                  const int32_t* __restrict__ segVoMap, int n_segs, int tm, int m, int k,
                  const float* __restrict__ B, float* __restrict__ C) {
  const int lane = threadIdx.x & 31;
  const int64_t seg = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (seg >= n_segs) return;
  const int32_t* rp = seg_rowPtr + seg * (tm + 1);
  for (int r = 0; r < tm; ++r) {
    const uint32_t tag = (uint32_t)__ldg(segVoMap + seg * tm + r);
    const int64_t row = (int64_t)(tag & 0x7fffffffu);
    if (row >= m) continue;  // padding rows of the last panel (tile.cu:93-96)
    const bool accumulate = (tag >> 31) != 0;
    const int beg = __ldg(rp + r), end = __ldg(rp + r + 1);
    for (int c0 = 0; c0 < k; c0 += 128) {
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      for (int z = beg; z < end; ++z) {
        const float2 cv = __ldg(reinterpret_cast<const float2*>(segNzCV) + z);
        const float* brow = B + (int64_t)(int)cv.x * k;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const int c = c0 + lane + 32 * t;
          if (c < k) acc[t] = fmaf(cv.y, __ldg(brow + c), acc[t]);
        }
      }
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int c = c0 + lane + 32 * t;
        if (c < k) {
          if (accumulate) atomicAdd(C + row * k + c, acc[t]);
          else C[row * k + c] = acc[t];
        }
      }
    }
  }
}

// Plain CSR (separate col / val arrays, as cuspmm receives them): warp per row.
__global__ void __launch_bounds__(128)
csr_rows_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                const float* __restrict__ val, int m, int k, const float* __restrict__ B,
                float* __restrict__ C) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= m) return;
  const int beg = __ldg(rowptr + row), end = __ldg(rowptr + row + 1);
  for (int c0 = 0; c0 < k; c0 += 128) {
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int z = beg; z < end; ++z) {
      const float a = __ldg(val + z);
      const float* brow = B + (int64_t)__ldg(col + z) * k;
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int c = c0 + lane + 32 * t;
        if (c < k) acc[t] = fmaf(a, __ldg(brow + c), acc[t]);
      }
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int c = c0 + lane + 32 * t;
      if (c < k) C[row * k + c] = acc[t];
    }
  }
}

static void report(const char* fn, int rc) {
  if (rc != GCNB_OK) fprintf(stderr, "gcn_b200: %s failed (%d): %s\n", fn, rc, gcnb_last_error());
}

static int launch_check(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error(GCNB_ERR_CUDA, "%s launch failed: %s", what, cudaGetErrorString(e));
  return GCNB_OK;
}

}  // namespace gcnb

using namespace gcnb;

extern "C" {

void flexspmm(int* seg_rowPtr, float* segNzCV, int* segVoMap, int* grouped_tailSeg, int* next_seg,
              int m, int n, int k, int n_segs, float* B, float* C) {
  (void)grouped_tailSeg; (void)next_seg; (void)n;  // no per-SM buckets in this engine
  int rc = require_gpu();
  if (rc == GCNB_OK && (!seg_rowPtr || !segNzCV || !segVoMap || !B || !C || m < 0 || k < 0 || n_segs < 0))
    rc = set_error(GCNB_ERR_ARG, "flexspmm: null pointer or negative size");
  if (rc == GCNB_OK && n_segs > 0 && k > 0) {
    const int tm = 8;  // fixed on both sides in the reference (flexspmm.cu:4, gcn6.py:340)
    const int64_t blocks = ((int64_t)n_segs + 3) / 4;
    flex_tiles_kernel<<<(unsigned)blocks, 128>>>(seg_rowPtr, segNzCV, segVoMap, n_segs, tm, m, k, B, C);
    rc = launch_check("flexspmm");
  }
  report("flexspmm", rc);
}

void cuspmm(float* rowPtr, int* col, float* vals, float* X, float* C, int m, int n, int nnz,
            int dim) {
  (void)n; (void)nnz;
  int rc = require_gpu();
  if (rc == GCNB_OK && (!rowPtr || !col || !vals || !X || !C || m < 0 || dim < 0))
    rc = set_error(GCNB_ERR_ARG, "cuspmm: null pointer or negative size");
  if (rc == GCNB_OK && m > 0 && dim > 0) {
    // the reference declares rowPtr as float* but passes int32 row offsets (cuspmm.cu:23,37)
    const int32_t* rp = reinterpret_cast<const int32_t*>(rowPtr);
    const int64_t blocks = ((int64_t)m + 3) / 4;
    csr_rows_kernel<<<(unsigned)blocks, 128>>>(rp, col, vals, m, dim, X, C);
    rc = launch_check("cuspmm");
  }
  report("cuspmm", rc);
}

void csr2tile(int* rowPtr, int* col, float* vals, int m, int n, int nnz, int* vo_mp,
              int* segVoMap, int* seg_rowPtr, float* segNzCV, int* grouped_tailSeg, int* next_seg,
              int tm, int* n_segs) {
  (void)vo_mp;
  int rc = require_gpu();  // the reference sizes its buckets from the current device (tile.cu:118-122)
  int n_sm = 0;
  if (rc == GCNB_OK) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
    n_sm = sm_count(dev);
    if (n_sm <= 0) rc = set_error(GCNB_ERR_CUDA, "csr2tile: cannot read the SM count");
  }
  if (rc == GCNB_OK && !n_segs) rc = set_error(GCNB_ERR_ARG, "csr2tile: n_segs is null");
  if (rc == GCNB_OK) {
    int64_t ns = 0;
    rc = gcnb_csr2tile(rowPtr, col, vals, m, n, nnz, tm, n_sm, segVoMap, seg_rowPtr, segNzCV,
                       grouped_tailSeg, next_seg, &ns);
    if (rc == GCNB_OK) n_segs[0] = (int)ns;
  }
  report("csr2tile", rc);
}

void permutate(float* B, int* voMp, int* labels, int m, int n, int k) {
  (void)labels; (void)m;  // label permutation is disabled in the reference (permutate.cu:17,35)
  int rc = require_gpu();
  if (rc == GCNB_OK && (!B || !voMp || n < 0 || k < 0)) rc = set_error(GCNB_ERR_ARG, "permutate: bad argument");
  if (rc == GCNB_OK && n > 0 && k > 0) {
    float* shadow = nullptr;
    const size_t bytes = (size_t)n * k * sizeof(float);
    cudaError_t e = cudaMallocAsync(&shadow, bytes, 0);
    if (e != cudaSuccess) rc = set_error(GCNB_ERR_ALLOC, "permutate: %s", cudaGetErrorString(e));
    if (rc == GCNB_OK) rc = gcnb_permute_rows(B, voMp, shadow, n, k, nullptr);
    if (rc == GCNB_OK) {
      e = cudaMemcpyAsync(B, shadow, bytes, cudaMemcpyDeviceToDevice, 0);
      if (e != cudaSuccess) rc = set_error(GCNB_ERR_CUDA, "permutate: %s", cudaGetErrorString(e));
    }
    if (shadow) cudaFreeAsync(shadow, 0);
  }
  report("permutate", rc);
}

}  // extern "C"
