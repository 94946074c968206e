// Exchange of feature rows between the GPUs of one box over NVLink peer memory.
//
// New relative to the reference (single GPU, no NCCL/MPI anywhere).  The row-partitioned SpMM needs,
// on every rank, the rows of the dense operand that its remote-column non-zeros touch (the "halo").
// Instead of all-gathering every block through NCCL, each rank publishes its block in an exchange
// buffer that the peers map through CUDA IPC, and pulls exactly its halo rows out of the owners'
// HBM with 128-bit P2P loads (gcnb_pull_rows): only rows that are needed cross NVLink, once each,
// and land compacted in the index space the remote block's column ids use.
#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include "common.h"

namespace gcnb {

constexpr int kMaxParts = 16;   // GPUs of one box (8 on a B200 HGX; 16 leaves room)
constexpr int kPullWarps = 8;
constexpr int kPullRowsPerWarp = 4;   // independent rows in flight per warp (NVLink latency ~2 us)

struct PeerBases {
  const char* p[kMaxParts];
};

// out[i, 0:row_bytes) = bases[part[i]][row[i] * src_pitch .. + row_bytes)
// Persistent, deliberately small grid (a couple of CTAs per SM, grid-stride over the halo): the pull
// is bound by NVLink, not by SM resources, and it runs at the same time as the local-block SpMM, which
// fills the register file; a grid sized by the halo would occupy every SM with warps that only wait on
// the link and serialise the two (measured on 2 GPUs, products-shaped: step 3.05 ms with a halo-sized
// grid, 2.75 / 2.69 / 2.45 ms with 4 / 2 / 1 CTAs per SM, profiles/r01_n_*; default 1).
// In flight: warps x kPullRowsPerWarp rows x row_bytes (148 x 8 x 4 x 400 B = 1.9 MB), about the
// ~2 us x 900 GB/s the link needs.
template <typename V>
__global__ void __launch_bounds__(kPullWarps * 32)
pull_rows_kernel(PeerBases bases, const int32_t* __restrict__ part, const int32_t* __restrict__ row,
                 char* __restrict__ out, int64_t src_pitch, int64_t out_pitch, int64_t n_rows,
                 int64_t first_row, int vecs_per_row) {
  const int lane = threadIdx.x & 31;
  const int64_t n_warps = (int64_t)gridDim.x * kPullWarps;
  for (int64_t w = (int64_t)blockIdx.x * kPullWarps + (threadIdx.x >> 5); w * kPullRowsPerWarp < n_rows;
       w += n_warps) {
    const int64_t r0 = w * kPullRowsPerWarp;
    const V* src[kPullRowsPerWarp];
    V* dst[kPullRowsPerWarp];
#pragma unroll
    for (int j = 0; j < kPullRowsPerWarp; ++j) {
      int64_t r = min(r0 + j, n_rows - 1);   // clamp: duplicates rewrite the same bytes
      r += first_row;                        // rotated start (see gcnb_pull_rows)
      if (r >= n_rows) r -= n_rows;
      src[j] = reinterpret_cast<const V*>(bases.p[__ldg(part + r)] + (int64_t)__ldg(row + r) * src_pitch);
      dst[j] = reinterpret_cast<V*>(out + r * out_pitch);
    }
    for (int i = lane; i < vecs_per_row; i += 32) {
      V v[kPullRowsPerWarp];
#pragma unroll
      for (int j = 0; j < kPullRowsPerWarp; ++j) v[j] = src[j][i];   // peer memory: plain (coherent) loads
#pragma unroll
      for (int j = 0; j < kPullRowsPerWarp; ++j) dst[j][i] = v[j];
    }
  }
}

}  // namespace gcnb

using namespace gcnb;

extern "C" {

int gcnb_peer_alloc(int device, int64_t bytes, void** ptr_out, unsigned char* handle_out) {
  GCNB_REQUIRE(ptr_out && handle_out, "null argument");
  GCNB_REQUIRE(bytes > 0, "bytes = %lld", (long long)bytes);
  *ptr_out = nullptr;
  int rc = require_gpu();
  if (rc != GCNB_OK) return rc;
  static_assert(sizeof(cudaIpcMemHandle_t) == GCNB_IPC_HANDLE_BYTES, "IPC handle size");
  DeviceGuard guard(device);
  GCNB_CHECK_CUDA(guard.status);
  void* p = nullptr;
  // plain cudaMalloc: IPC handles cannot be taken from stream-ordered (pool) allocations
  GCNB_CHECK_CUDA(cudaMalloc(&p, (size_t)bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    return set_error(GCNB_ERR_CUDA, "cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
  }
  memcpy(handle_out, &h, sizeof(h));
  *ptr_out = p;
  return GCNB_OK;
}

int gcnb_peer_free(int device, void* ptr) {
  if (!ptr) return GCNB_OK;
  DeviceGuard guard(device);
  GCNB_CHECK_CUDA(guard.status);
  GCNB_CHECK_CUDA(cudaFree(ptr));
  return GCNB_OK;
}

int gcnb_peer_open(int device, const unsigned char* handle, void** ptr_out) {
  GCNB_REQUIRE(handle && ptr_out, "null argument");
  *ptr_out = nullptr;
  int rc = require_gpu();
  if (rc != GCNB_OK) return rc;
  DeviceGuard guard(device);
  GCNB_CHECK_CUDA(guard.status);
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  void* p = nullptr;
  // maps the owner's allocation into this process and enables peer access to its device
  GCNB_CHECK_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *ptr_out = p;
  return GCNB_OK;
}

int gcnb_peer_close(int device, void* ptr) {
  if (!ptr) return GCNB_OK;
  DeviceGuard guard(device);
  GCNB_CHECK_CUDA(guard.status);
  GCNB_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
  return GCNB_OK;
}

int gcnb_pull_rows(const void* const* bases, int n_parts, int64_t src_pitch_bytes,
                   const int32_t* part, const int32_t* row, void* out, int64_t out_pitch_bytes,
                   int64_t n_rows, int64_t row_bytes, int64_t first_row, void* stream) {
  GCNB_REQUIRE(n_rows >= 0 && row_bytes >= 0, "negative size");
  GCNB_REQUIRE(first_row >= 0 && (first_row < n_rows || n_rows == 0), "first_row out of range");
  if (n_rows == 0 || row_bytes == 0) return GCNB_OK;
  GCNB_REQUIRE(bases && part && row && out, "null pointer");
  GCNB_REQUIRE(n_parts >= 1 && n_parts <= kMaxParts, "n_parts = %d (1..%d)", n_parts, kMaxParts);
  GCNB_REQUIRE(src_pitch_bytes >= row_bytes && out_pitch_bytes >= row_bytes, "pitch smaller than a row");
  int rc = require_gpu();
  if (rc != GCNB_OK) return rc;
  const int dev = pointer_device(out);
  GCNB_REQUIRE(dev >= 0, "out is not a device pointer");
  DeviceGuard guard(dev);
  GCNB_CHECK_CUDA(guard.status);
  PeerBases pb;
  uintptr_t align = (uintptr_t)out | (uintptr_t)src_pitch_bytes | (uintptr_t)out_pitch_bytes | (uintptr_t)row_bytes;
  for (int q = 0; q < kMaxParts; ++q) {
    pb.p[q] = q < n_parts ? (const char*)bases[q] : nullptr;
    align |= (uintptr_t)pb.p[q];   // a part this rank never reads may be null
  }
  const int64_t warps = (n_rows + kPullRowsPerWarp - 1) / kPullRowsPerWarp;
  static const int ctas_per_sm = getenv("GCNB_PULL_CTAS") ? atoi(getenv("GCNB_PULL_CTAS")) : 1;
  const int64_t blocks = std::min<int64_t>((warps + kPullWarps - 1) / kPullWarps,
                                           (int64_t)std::max(1, sm_count(dev)) * std::max(1, ctas_per_sm));
  cudaStream_t s = (cudaStream_t)stream;
  if (align % 16 == 0)
    pull_rows_kernel<uint4><<<(unsigned)blocks, kPullWarps * 32, 0, s>>>(
        pb, part, row, (char*)out, src_pitch_bytes, out_pitch_bytes, n_rows, first_row, (int)(row_bytes / 16));
  else if (align % 4 == 0)
    pull_rows_kernel<uint32_t><<<(unsigned)blocks, kPullWarps * 32, 0, s>>>(
        pb, part, row, (char*)out, src_pitch_bytes, out_pitch_bytes, n_rows, first_row, (int)(row_bytes / 4));
  else
    pull_rows_kernel<unsigned char><<<(unsigned)blocks, kPullWarps * 32, 0, s>>>(
        pb, part, row, (char*)out, src_pitch_bytes, out_pitch_bytes, n_rows, first_row, (int)row_bytes);
  GCNB_CHECK_CUDA(cudaGetLastError());
  return GCNB_OK;
}

}  // extern "C"
