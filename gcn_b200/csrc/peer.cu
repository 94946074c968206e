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
// (An L2 evict_first priority on the halo rows written here and on the peer loads changed nothing beside the local
// block, profiles/r02_ax_timeline_n4.jsonl; the accesses are plain.)
// Register budget: the kernel runs beside the local-block SpMM, whose occupancy is set by the register file.  At 80
// registers x 256 threads the pull held 31 % of it and the local block lost a third of its warps (4 GPUs: 1.05 ms
// beside the pull against 0.77 alone, profiles/r02_aw_timeline_n4.jsonl).  Hence: the peer bases are read from the
// constant bank (__grid_constant__; a dynamically indexed by-value array would be copied to the stack), row indices
// stay 32-bit until they are used, the vector loop is not unrolled (a 400-byte row is one trip), and the launch
// bounds cap the kernel at 51 registers (13 k of the SM's 64 k: the fp32 SpMM keeps 7 of its 9 CTAs; 40 registers
// would not give it an eighth and spill here).
template <typename V>
__global__ void __launch_bounds__(kPullWarps * 32, 5)
pull_rows_kernel(const __grid_constant__ PeerBases bases, const int32_t* __restrict__ part,
                 const int32_t* __restrict__ row, char* __restrict__ out, int64_t src_pitch, int64_t out_pitch,
                 int64_t n_rows, int64_t first_row, int vecs_per_row) {
  const int lane = threadIdx.x & 31;
  const int64_t n_warps = (int64_t)gridDim.x * kPullWarps;
#pragma unroll 1
  for (int64_t w = (int64_t)blockIdx.x * kPullWarps + (threadIdx.x >> 5); w * kPullRowsPerWarp < n_rows;
       w += n_warps) {
    const int64_t r0 = w * kPullRowsPerWarp;
    const V* src[kPullRowsPerWarp];
    int32_t rr[kPullRowsPerWarp];   // halo positions fit 32 bits (checked by gcnb_pull_rows)
#pragma unroll
    for (int j = 0; j < kPullRowsPerWarp; ++j) {
      int64_t r = min(r0 + j, n_rows - 1);   // clamp: duplicates rewrite the same bytes
      r += first_row;                        // rotated start (see gcnb_pull_rows)
      if (r >= n_rows) r -= n_rows;
      rr[j] = (int32_t)r;
      src[j] = reinterpret_cast<const V*>(bases.p[__ldg(part + r)] + (int64_t)__ldg(row + r) * src_pitch);
    }
#pragma unroll 1
    for (int i = lane; i < vecs_per_row; i += 32) {
      V v[kPullRowsPerWarp];
#pragma unroll
      for (int j = 0; j < kPullRowsPerWarp; ++j) v[j] = src[j][i];   // peer memory: plain (coherent) loads
#pragma unroll
      for (int j = 0; j < kPullRowsPerWarp; ++j) *(reinterpret_cast<V*>(out + (int64_t)rr[j] * out_pitch) + i) = v[j];
    }
  }
}

}  // namespace gcnb

namespace gcnb {

// ---- the same pull through the TMA unit: no registers, no warps waiting on the link ---------------------
// The LDG kernel above keeps warps x 4 rows in flight and holds them in registers (1.9 MB over the chip,
// 540 GB/s inbound at 8 GPUs = 60 % of a link direction), and its CTAs take register file from the local-block
// SpMM that runs beside it.  Here ONE thread per warp drives a ring of D row-sized bulk copies
// (cp.async.bulk: peer HBM -> shared memory, completion on an mbarrier; shared memory -> local HBM as a bulk
// group): L = 3D/4 loads in flight per producer thread, 4 producer threads per CTA, 148 CTAs — 5.7 MB at
// D = 32 and 400-byte rows — with ~20 registers per thread and D x row bytes of shared memory per warp.
// Rows must be multiples of 16 bytes at 16-byte aligned addresses (fp32 rows with k % 4 == 0).
constexpr int kTmaPullWarps = 4;

__device__ __forceinline__ void pull_wait_read(int pending) {
  // cp.async.bulk.wait_group.read takes an immediate
  switch (pending) {
    case 0: asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); break;
    case 1: asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); break;
    case 3: asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory"); break;
    default: asm volatile("cp.async.bulk.wait_group.read 7;" ::: "memory"); break;
  }
}

template <int D>
__global__ void __launch_bounds__(kTmaPullWarps * 32)
pull_rows_tma_kernel(PeerBases bases, const int32_t* __restrict__ part, const int32_t* __restrict__ row,
                     char* __restrict__ out, int64_t src_pitch, int64_t out_pitch, int64_t n_rows,
                     int64_t first_row, uint32_t row_bytes) {
  constexpr int L = D * 3 / 4;          // loads in flight per producer thread
  constexpr int PEND = D - 1 - L;       // stores that may still be reading shared memory when a slot is re-used
  static_assert(PEND == 0 || PEND == 1 || PEND == 3 || PEND == 7, "wait_group immediates");
  extern __shared__ __align__(128) unsigned char pull_smem[];
  const int warp = threadIdx.x >> 5;
  if ((threadIdx.x & 31) != 0) return;  // one thread per warp drives the copies; nothing below synchronises
  const uint32_t smem0 = (uint32_t)__cvta_generic_to_shared(pull_smem);
  const uint32_t slots = smem0 + (uint32_t)warp * D * row_bytes;
  const uint32_t bars = smem0 + (uint32_t)kTmaPullWarps * D * row_bytes + (uint32_t)warp * D * 8u;
#pragma unroll 1
  for (int d = 0; d < D; ++d) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bars + 8u * d));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  const int64_t g = (int64_t)blockIdx.x * kTmaPullWarps + warp, G = (int64_t)gridDim.x * kTmaPullWarps;
  const int64_t cnt = g < n_rows ? (n_rows - g + G - 1) / G : 0;
  auto rotated = [&](int64_t t) {
    int64_t r = g + t * G + first_row;
    return r >= n_rows ? r - n_rows : r;
  };
#pragma unroll 1
  for (int64_t t = 0; t < cnt + L; ++t) {
    if (t < cnt) {
      const uint32_t d = (uint32_t)(t % D);
      if (t >= D) pull_wait_read(PEND);                         // the store that last used this slot has read it
      const int64_t r = rotated(t);
      const char* src = bases.p[__ldg(part + r)] + (int64_t)__ldg(row + r) * src_pitch;
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bars + 8u * d), "r"(row_bytes) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(slots + d * row_bytes), "l"(src), "r"(row_bytes), "r"(bars + 8u * d) : "memory");
    }
    const int64_t u = t - L;
    if (u >= 0 && u < cnt) {
      const uint32_t d = (uint32_t)(u % D);
      const uint32_t parity = (uint32_t)((u / D) & 1);
      uint32_t done = 0;
      while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bars + 8u * d), "r"(parity) : "memory");
      char* dst = out + rotated(u) * out_pitch;
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                   ::"l"(dst), "r"(slots + d * row_bytes), "r"(row_bytes) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// ---- rendezvous of the ranks through flags in peer memory ----------------------------------------------
// flags[q] of rank r's array is written by rank q only.  Thread q of the one CTA: store this call's epoch into
// OUR word of rank q's array (release, system scope: everything this GPU wrote before — the operand block in the
// exchange slot — is visible to a peer that has seen the flag), then wait until rank q's word in our own array
// has reached the epoch.  Epochs only grow; the comparison is on the signed difference, so a wrap is harmless.
// A peer that never arrives would leave the kernel spinning: after ~10 s it traps (a failed launch is reported
// by the next CUDA call; a hung GPU is not).
__global__ void peer_barrier_kernel(PeerBases flags, int n_parts, int rank, uint32_t epoch) {
  const int q = threadIdx.x;
  if (q >= n_parts || q == rank) return;
  uint32_t* theirs = reinterpret_cast<uint32_t*>(const_cast<char*>(flags.p[q])) + rank;
  const uint32_t* ours = reinterpret_cast<const uint32_t*>(flags.p[rank]) + q;
  asm volatile("fence.acq_rel.sys;" ::: "memory");
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(theirs), "r"(epoch) : "memory");
  unsigned long long t0 = 0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (unsigned spins = 0;; ++spins) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(ours) : "memory");
    if ((int32_t)(v - epoch) >= 0) break;
    if ((spins & 1023u) == 1023u) {
      unsigned long long t1;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > 10000000000ull) __trap();
    }
  }
}

template <int D>
static cudaError_t launch_pull_tma(const PeerBases& pb, const int32_t* part, const int32_t* row, char* out,
                                   int64_t src_pitch, int64_t out_pitch, int64_t n_rows, int64_t first_row,
                                   uint32_t row_bytes, int blocks, cudaStream_t s) {
  const size_t smem = (size_t)kTmaPullWarps * D * row_bytes + (size_t)kTmaPullWarps * D * 8;
  cudaError_t e = cudaFuncSetAttribute(pull_rows_tma_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  pull_rows_tma_kernel<D><<<blocks, kTmaPullWarps * 32, smem, s>>>(pb, part, row, out, src_pitch, out_pitch, n_rows,
                                                                     first_row, row_bytes);
  return cudaGetLastError();
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
  GCNB_REQUIRE(n_rows < 0x7fffffffLL, "n_rows = %lld does not fit 32 bits", (long long)n_rows);
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
  // GCNB_PULL_TMA: 0 = the LDG kernel, D = 8 / 16 / 32 = ring depth of the bulk-copy kernel (default 32, halved
  // until 4 warps x D rows fit 96 KB of shared memory)
  static const int pull_tma = getenv("GCNB_PULL_TMA") ? atoi(getenv("GCNB_PULL_TMA")) : 0;
  if (pull_tma > 0 && align % 16 == 0 && row_bytes <= 3072) {
    int D = pull_tma >= 32 ? 32 : (pull_tma >= 16 ? 16 : 8);
    while (D > 8 && (int64_t)kTmaPullWarps * D * row_bytes > 96 * 1024) D /= 2;
    const int64_t producers = std::min<int64_t>(n_rows, (int64_t)std::max(1, sm_count(dev)) * std::max(1, ctas_per_sm) * kTmaPullWarps);
    const int tb = (int)((producers + kTmaPullWarps - 1) / kTmaPullWarps);
    cudaError_t e;
    if (D == 32) e = launch_pull_tma<32>(pb, part, row, (char*)out, src_pitch_bytes, out_pitch_bytes, n_rows, first_row, (uint32_t)row_bytes, tb, s);
    else if (D == 16) e = launch_pull_tma<16>(pb, part, row, (char*)out, src_pitch_bytes, out_pitch_bytes, n_rows, first_row, (uint32_t)row_bytes, tb, s);
    else e = launch_pull_tma<8>(pb, part, row, (char*)out, src_pitch_bytes, out_pitch_bytes, n_rows, first_row, (uint32_t)row_bytes, tb, s);
    GCNB_CHECK_CUDA(e);
    return GCNB_OK;
  }
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

int gcnb_peer_barrier(const void* const* flag_bases, int n_parts, int rank, uint32_t epoch, void* stream) {
  GCNB_REQUIRE(flag_bases != nullptr, "flag_bases is null");
  GCNB_REQUIRE(n_parts >= 1 && n_parts <= kMaxParts, "n_parts = %d (1..%d)", n_parts, kMaxParts);
  GCNB_REQUIRE(rank >= 0 && rank < n_parts, "rank %d outside 0..%d", rank, n_parts - 1);
  if (n_parts == 1) return GCNB_OK;
  int rc = require_gpu();
  if (rc != GCNB_OK) return rc;
  PeerBases pb;
  for (int q = 0; q < kMaxParts; ++q) {
    pb.p[q] = q < n_parts ? (const char*)flag_bases[q] : nullptr;
    GCNB_REQUIRE(q >= n_parts || (pb.p[q] != nullptr && (uintptr_t)pb.p[q] % 4 == 0), "flag array %d null or unaligned", q);
  }
  const int dev = pointer_device(flag_bases[rank]);
  GCNB_REQUIRE(dev >= 0, "flag array of this rank is not a device pointer");
  DeviceGuard guard(dev);
  GCNB_CHECK_CUDA(guard.status);
  peer_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(pb, n_parts, rank, epoch);
  GCNB_CHECK_CUDA(cudaGetLastError());
  return GCNB_OK;
}

}  // extern "C"
