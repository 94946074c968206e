// Does the Blackwell tensor-map TMA gather (cp.async.bulk.tensor.2d ... tile::gather4: four rows of a
// 2-D tensor per instruction, written to shared memory by the async proxy, completion on an mbarrier)
// pull feature rows out of L2 / HBM faster than per-lane LDG.128?  
//
//   tma_gather4_probe <table_mb> <row_bytes> <consumer_warps> <ctas_per_sm> <stages> [window] [pitch]
//
// Kernel shape (what an SpMM built on it would look like): one producer warp per CTA walks the index
// stream 32 rows at a time (one coalesced index load; lanes 0..7 each issue ONE gather4 = 4 rows), a ring
// of `stages` shared-memory stages of 32 rows each, full/empty mbarriers per stage; the consumer warps
// read whole rows back with LDS.128 (lane l owns bytes [16 l, 16 l + 16) of the row, as the SpMM kernels
// do) and accumulate.  Rows in flight per SM = stages x 32 x ctas_per_sm, limited by shared memory only
// (no registers are held by a gather in flight).
// The LDG form of the same loop over the same index stream is timed next to it: (consumer_warps + 1)
// warps per CTA, U = 8 independent gathers per warp.
// Output: one JSON line per mode.  Every wait is bounded (trap after ~2^28 polls), so a bad tensor map
// faults instead of hanging the box.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("{\"error\": \"%s at line %d\"}\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  uint32_t spins = 0;
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (++spins > (1u << 28)) __trap();
  }
}

constexpr int kStageRows = 32;

__global__ void __launch_bounds__(1024, 1)
k_g4(const __grid_constant__ CUtensorMap tmap, const int* __restrict__ idx, int n_per_cta, int row_bytes,
     int op_bytes, int stages, float* out) {
  extern __shared__ __align__(128) char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, cw = (blockDim.x >> 5) - 1;
  const int stage_bytes = (kStageRows / 4) * op_bytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)stages * stage_bytes);
  uint64_t* empty = full + stages;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(full + s)));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(empty + s)), "r"(cw));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int* my = idx + (size_t)blockIdx.x * n_per_cta;
  const int T = n_per_cta / kStageRows;
  if (warp == 0) {
    // producer
    int r = __ldg(my + lane);
    for (int t = 0; t < T; ++t) {
      const int s = t % stages;
      mbar_wait(s32(empty + s), ((t / stages) & 1) ^ 1);
      const int r0 = __shfl_sync(0xffffffffu, r, (4 * lane) & 31), r1 = __shfl_sync(0xffffffffu, r, (4 * lane + 1) & 31);
      const int r2 = __shfl_sync(0xffffffffu, r, (4 * lane + 2) & 31), r3 = __shfl_sync(0xffffffffu, r, (4 * lane + 3) & 31);
      if (t + 1 < T) r = __ldg(my + (t + 1) * kStageRows + lane);
      const uint32_t bar = s32(full + s);
      if (lane == 0)
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(kStageRows * row_bytes) : "memory");
      __syncwarp();
      if (lane < kStageRows / 4) {
        const uint32_t dst = s32(smem + (size_t)s * stage_bytes + (size_t)lane * op_bytes);
        asm volatile(
            "cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
            ::"r"(dst), "l"(&tmap), "r"(0), "r"(r0), "r"(r1), "r"(r2), "r"(r3), "r"(bar) : "memory");
      }
    }
  } else {
    const int c = warp - 1;
    float4 acc = make_float4(0, 0, 0, 0);
    const bool act = lane * 16 < row_bytes;
    for (int t = 0; t < T; ++t) {
      const int s = t % stages;
      mbar_wait(s32(full + s), (t / stages) & 1);
      const char* base = smem + (size_t)s * stage_bytes;
      for (int r = c; r < kStageRows; r += cw) {
        if (act) {
          const float4 v = *reinterpret_cast<const float4*>(base + (r >> 2) * op_bytes + (r & 3) * row_bytes + lane * 16);
          acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
      }
      __syncwarp();
      if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(empty + s)) : "memory");
    }
    atomicAdd(out + blockIdx.x, acc.x + acc.y + acc.z + acc.w);   // small integers: exact in fp32
  }
}

__global__ void fill_kernel(float* t, size_t rows, int pitch_f, int row_f) {
  for (size_t r = blockIdx.x; r < rows; r += gridDim.x)
    for (int j = threadIdx.x; j < row_f; j += blockDim.x) t[r * pitch_f + j] = (float)(r % 3);
}

template <int U>
__global__ void k_ldg(const char* __restrict__ table, const int* __restrict__ idx, int n_per_cta, int row_bytes,
                      int pitch, float* out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int* my = idx + (size_t)blockIdx.x * n_per_cta;
  const bool act = lane * 16 < row_bytes;
  const char* base = table + (act ? lane * 16 : 0);
  float4 acc = make_float4(0, 0, 0, 0);
  // warp w takes groups of U rows round-robin
  for (int i = warp * U; i + U <= n_per_cta; i += nw * U) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = __ldg(reinterpret_cast<const float4*>(base + (size_t)__ldg(my + i + u) * pitch));
#pragma unroll
    for (int u = 0; u < U; ++u) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
  }
  if (act) atomicAdd(out + blockIdx.x, acc.x + acc.y + acc.z + acc.w);
}

typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
  const int table_mb = argc > 1 ? atoi(argv[1]) : 64;
  const int row_bytes = argc > 2 ? atoi(argv[2]) : 400;
  const int cw = argc > 3 ? atoi(argv[3]) : 8;
  const int cps = argc > 4 ? atoi(argv[4]) : 1;
  int stages = argc > 5 ? atoi(argv[5]) : 0;
  const int window = argc > 6 ? atoi(argv[6]) : 0;
  const int pitch = argc > 7 && atoi(argv[7]) > 0 ? atoi(argv[7]) : row_bytes;
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  const int sms = prop.multiProcessorCount;
  const size_t rows = (size_t)table_mb * 1024 * 1024 / pitch;
  const int op_bytes = (4 * row_bytes + 127) & ~127;
  const int stage_bytes = (kStageRows / 4) * op_bytes;
  if (stages <= 0) stages = (int)((200 * 1024 / cps - 1024) / stage_bytes);
  const size_t smem = (size_t)stages * stage_bytes + (size_t)stages * 16 + 128;
  const int n_per_cta = 32768 / cps;   // rows gathered per CTA
  const int ctas = sms * cps * 2;      // two waves
  std::vector<int> h((size_t)ctas * n_per_cta);
  uint64_t s = 88172645463325252ull;
  size_t w = 0;
  for (int c = 0; c < ctas; ++c) {
    const size_t center = (size_t)((double)c / ctas * rows);
    for (int i = 0; i < n_per_cta; ++i) {
      s ^= s << 13; s ^= s >> 7; s ^= s << 17;
      if (window > 0) { long long r = (long long)center + (long long)((double)i / n_per_cta * rows / ctas) + (long long)(s % (2 * window)) - window; r = ((r % (long long)rows) + rows) % rows; h[w++] = (int)r; }
      else h[w++] = (int)(s % rows);
    }
  }
  char* table; int* idx; float* out;
  CK(cudaMalloc(&table, rows * pitch + 1024)); CK(cudaMemset(table, 0, rows * pitch + 1024));
  CK(cudaMalloc(&idx, h.size() * 4)); CK(cudaMemcpy(idx, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&out, ctas * 4));
  fill_kernel<<<sms * 8, 128>>>((float*)table, rows, pitch / 4, row_bytes / 4); CK(cudaDeviceSynchronize());
  std::vector<float> want(ctas), got(ctas);
  for (int c = 0; c < ctas; ++c) {
    double e = 0;
    for (int i = 0; i < n_per_cta; ++i) e += h[(size_t)c * n_per_cta + i] % 3;
    want[c] = (float)(e * (row_bytes / 4));
  }
  auto check = [&](const char* mode) {
    CK(cudaMemcpy(got.data(), out, ctas * 4, cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int c = 0; c < ctas; ++c) bad += got[c] != want[c];
    if (bad) printf("{\"error\": \"%s: %d of %d CTA checksums differ (first: got %.1f want %.1f)\"}\n", mode, bad, ctas, got[0], want[0]);
    return bad == 0;
  };
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const double total_rows = (double)ctas * n_per_cta;
  const double clk = prop.clockRate * 1e3;
  auto report = [&](const char* mode, int depth, int warps, float ms) {
    printf("{\"mode\": \"%s\", \"depth\": %d, \"table_mb\": %d, \"row_bytes\": %d, \"pitch\": %d, \"warps_per_cta\": %d, \"ctas_per_sm\": %d, "
           "\"window\": %d, \"rows_in_flight_per_sm\": %d, \"ms\": %.4f, \"gather_GBps\": %.1f, \"cyc_per_row_per_sm\": %.2f}\n",
           mode, depth, table_mb, row_bytes, pitch, warps, cps, window,
           mode[0] == 'g' ? depth * kStageRows * cps : depth * warps * cps, ms, total_rows * row_bytes / ms / 1e6,
           ms * 1e-3 * clk / (total_rows / sms));
    fflush(stdout);
  };
  // LDG reference
  {
    const int nw = cw + 1;
    CK(cudaMemset(out, 0, ctas * 4));
    k_ldg<8><<<ctas, nw * 32>>>(table, idx, n_per_cta, row_bytes, pitch, out); CK(cudaDeviceSynchronize());
    check("ldg");
    CK(cudaEventRecord(e0));
    for (int i = 0; i < 3; ++i) k_ldg<8><<<ctas, nw * 32>>>(table, idx, n_per_cta, row_bytes, pitch, out);
    CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); report("ldg", 8, nw, ms / 3);
  }
  // tensor-map gather4
  EncodeTiled encode = nullptr;
  cudaDriverEntryPointQueryResult qres;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&encode, cudaEnableDefault, &qres));
  if (!encode || qres != cudaDriverEntryPointSuccess) { printf("{\"error\": \"cuTensorMapEncodeTiled unavailable\"}\n"); return 1; }
  CUtensorMap tmap;
  const cuuint64_t gdim[2] = {(cuuint64_t)(row_bytes / 4), (cuuint64_t)rows};
  const cuuint64_t gstride[1] = {(cuuint64_t)pitch};
  const cuuint32_t box[2] = {(cuuint32_t)(row_bytes / 4), 1};
  const cuuint32_t estr[2] = {1, 1};
  CUresult rc = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, table, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (rc != CUDA_SUCCESS) { printf("{\"error\": \"cuTensorMapEncodeTiled rc=%d\"}\n", (int)rc); return 1; }
  CK(cudaFuncSetAttribute(k_g4, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaMemset(out, 0, ctas * 4));
  k_g4<<<ctas, (cw + 1) * 32, smem>>>(tmap, idx, n_per_cta, row_bytes, op_bytes, stages, out);
  CK(cudaDeviceSynchronize());
  if (!check("gather4")) return 1;
  CK(cudaEventRecord(e0));
  for (int i = 0; i < 3; ++i) k_g4<<<ctas, (cw + 1) * 32, smem>>>(tmap, idx, n_per_cta, row_bytes, op_bytes, stages, out);
  CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
  float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); report("gather4", stages, cw + 1, ms / 3);
  return 0;
}
