// Can bulk async copies (cp.async.bulk, the non-tensor TMA path) gather feature rows faster than
// per-lane LDG.128?  Each warp owns a private shared-memory ring of S row slots with one mbarrier
// per slot; lane 0 issues one bulk copy per gathered row, all lanes wait on the slot's mbarrier,
// read the row back with LDS.128 and accumulate.  Reported next to the LDG form of the same loop.
//   tma_gather_probe <table_mb> <row_bytes> <warps_per_cta> <ctas_per_sm> <window>
// Output: one JSON line per (mode, ring depth).
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int S>
__global__ void k_bulk(const char* __restrict__ table, const int* __restrict__ idx, int n, int row_bytes,
                       int pitch, float* out) {
  extern __shared__ __align__(128) char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int slot_bytes = (row_bytes + 127) & ~127;
  char* ring = smem + (size_t)warp * S * slot_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)nw * S * slot_bytes) + warp * S;
  const int* my = idx + ((size_t)blockIdx.x * nw + warp) * n;
  if (lane == 0) {
    for (int s = 0; s < S; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(bars + s)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  auto issue = [&](int i) {   // lane 0 only
    const int s = i % S;
    const uint32_t bar = s32(bars + s);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(row_bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(s32(ring + (size_t)s * slot_bytes)), "l"(table + (size_t)__ldg(my + i) * pitch),
                   "r"(row_bytes), "r"(bar) : "memory");
  };
  if (lane == 0)
    for (int i = 0; i < S && i < n; ++i) issue(i);
  float4 acc = make_float4(0, 0, 0, 0);
  const bool act = lane * 16 < row_bytes;
  for (int i = 0; i < n; ++i) {
    const int s = i % S;
    const uint32_t bar = s32(bars + s), parity = (i / S) & 1;
    uint32_t done = 0;
    while (!done)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (act) {
      const float4 v = *reinterpret_cast<const float4*>(ring + (size_t)s * slot_bytes + lane * 16);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    __syncwarp();   // every lane has read the slot before it is refilled
    if (lane == 0 && i + S < n) issue(i + S);
  }
  if (acc.x + acc.y + acc.z + acc.w == 12345.678f) out[0] = acc.x;
}

template <int U>
__global__ void k_ldg(const char* __restrict__ table, const int* __restrict__ idx, int n, int row_bytes,
                      int pitch, float* out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int* my = idx + ((size_t)blockIdx.x * nw + warp) * n;
  const bool act = lane * 16 < row_bytes;
  const char* base = table + (act ? lane * 16 : 0);
  float4 acc = make_float4(0, 0, 0, 0);
  for (int i = 0; i < n; i += U) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = __ldg(reinterpret_cast<const float4*>(base + (size_t)__ldg(my + i + u) * pitch));
#pragma unroll
    for (int u = 0; u < U; ++u) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
  }
  if (act && acc.x + acc.y + acc.z + acc.w == 12345.678f) out[0] = acc.x;
}

int main(int argc, char** argv) {
  const int table_mb = argc > 1 ? atoi(argv[1]) : 64;
  const int row_bytes = argc > 2 ? atoi(argv[2]) : 400;
  const int wpc = argc > 3 ? atoi(argv[3]) : 8;
  const int cps = argc > 4 ? atoi(argv[4]) : 4;
  const int window = argc > 5 ? atoi(argv[5]) : 0;
  const int pitch = argc > 6 ? atoi(argv[6]) : row_bytes;
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  const int sms = prop.multiProcessorCount;
  const size_t rows = (size_t)table_mb * 1024 * 1024 / pitch;
  const int n = 1024;
  const int ctas = sms * cps * 4;   // 4 waves
  std::vector<int> h((size_t)ctas * wpc * n);
  uint64_t s = 88172645463325252ull;
  size_t w = 0;
  for (size_t wi = 0; wi < (size_t)ctas * wpc; ++wi) {
    const size_t center = (size_t)((double)wi / (ctas * wpc) * rows);
    for (int i = 0; i < n; ++i) {
      s ^= s << 13; s ^= s >> 7; s ^= s << 17;
      if (window > 0) { long long r = (long long)center + (long long)(s % (2 * window)) - window; r = ((r % (long long)rows) + rows) % rows; h[w++] = (int)r; }
      else h[w++] = (int)(s % rows);
    }
  }
  char* table; int* idx; float* out;
  CK(cudaMalloc(&table, rows * pitch + 1024)); CK(cudaMemset(table, 0, rows * pitch + 1024));
  CK(cudaMalloc(&idx, h.size() * 4)); CK(cudaMemcpy(idx, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&out, 64));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const double total_rows = (double)ctas * wpc * n;
  auto report = [&](const char* mode, int depth, float ms) {
    printf("{\"mode\": \"%s\", \"depth\": %d, \"table_mb\": %d, \"row_bytes\": %d, \"pitch\": %d, \"warps_per_cta\": %d, \"ctas_per_sm\": %d, "
           "\"window\": %d, \"ms\": %.4f, \"gather_GBps\": %.1f, \"cyc_per_row_per_sm\": %.2f}\n",
           mode, depth, table_mb, row_bytes, pitch, wpc, cps, window, ms, total_rows * row_bytes / ms / 1e6,
           ms * 1e-3 * 1.965e9 / (total_rows / sms));
  };
#define RUN_BULK(S) { const int slot = (row_bytes + 127) & ~127; const size_t sm = (size_t)wpc * S * slot + (size_t)wpc * S * 8 + 128; \
    if (sm * cps <= 220 * 1024) { CK(cudaFuncSetAttribute(k_bulk<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm)); \
    k_bulk<S><<<ctas, wpc * 32, sm>>>(table, idx, n, row_bytes, pitch, out); CK(cudaDeviceSynchronize()); \
    CK(cudaEventRecord(e0)); k_bulk<S><<<ctas, wpc * 32, sm>>>(table, idx, n, row_bytes, pitch, out); CK(cudaEventRecord(e1)); \
    CK(cudaDeviceSynchronize()); float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); report("bulk", S, ms); } }
#define RUN_LDG(U) { k_ldg<U><<<ctas, wpc * 32>>>(table, idx, n, row_bytes, pitch, out); CK(cudaDeviceSynchronize()); \
    CK(cudaEventRecord(e0)); k_ldg<U><<<ctas, wpc * 32>>>(table, idx, n, row_bytes, pitch, out); CK(cudaEventRecord(e1)); \
    CK(cudaDeviceSynchronize()); float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); report("ldg", U, ms); }
  RUN_LDG(4) RUN_LDG(8)
  RUN_BULK(2) RUN_BULK(4) RUN_BULK(8) RUN_BULK(16)
  return 0;
}
