// Does one warp overlap several independent gathered row loads?  (follow-up to gather_probe: with one
// warp per SM the time per batch of U loads grew linearly with U.)
//   mlp_probe <row_bytes> <warps_per_sm>
// MODE 0: __ldg float4   1: asm volatile ld.global.nc.v4.f32   2: asm volatile ld.global.v4.f32 (coherent path)
// Row indices come from shared memory (no shuffle / no dependent global load in the timed loop).
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

template <int MODE>
__device__ __forceinline__ float4 ld(const char* p) {
  float4 v;
  if (MODE == 0) v = __ldg(reinterpret_cast<const float4*>(p));
  else if (MODE == 1) asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  else asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

template <int U, int MODE>
__global__ void k(const char* __restrict__ table, const int* __restrict__ idx, int n, int row_bytes, float* out, long long* cyc) {
  extern __shared__ int s_idx[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int i = threadIdx.x; i < n * nw; i += blockDim.x) s_idx[i] = idx[(size_t)blockIdx.x * n * nw + i];
  __syncthreads();
  const int* my = s_idx + warp * n;
  const bool act = lane * 16 < row_bytes;
  const char* base = table + (act ? lane * 16 : 0);
  float4 acc = make_float4(0, 0, 0, 0);
  const long long t0 = clock64();
  for (int i = 0; i < n; i += U) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = ld<MODE>(base + (size_t)my[i + u] * row_bytes);
#pragma unroll
    for (int u = 0; u < U; ++u) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
  }
  const long long t1 = clock64();
  if (acc.x + acc.y + acc.z + acc.w == 12345.678f) out[0] = acc.x;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

static int g_table_mb = 64;
template <int U, int MODE>
void run(const char* table, const int* idx, int n, int row_bytes, float* out, long long* cyc, int sms, int wps, const char* nm) {
  k<U, MODE><<<sms, wps * 32, n * wps * 4>>>(table, idx, n, row_bytes, out, cyc); CK(cudaDeviceSynchronize());
  k<U, MODE><<<sms, wps * 32, n * wps * 4>>>(table, idx, n, row_bytes, out, cyc); CK(cudaDeviceSynchronize());
  std::vector<long long> h(sms); CK(cudaMemcpy(h.data(), cyc, sms * 8, cudaMemcpyDeviceToHost));
  double s = 0; for (auto c : h) s += c; s /= sms;
  printf("{\"mode\": \"%s\", \"table_mb\": %d, \"row_bytes\": %d, \"warps_per_sm\": %d, \"U\": %d, \"cycles_per_row_per_warp\": %.1f, \"cycles_per_batch\": %.1f}\n",
         nm, g_table_mb, row_bytes, wps, U, s / n, s / n * U);
}

int main(int argc, char** argv) {
  const int row_bytes = argc > 1 ? atoi(argv[1]) : 400;
  const int wps = argc > 2 ? atoi(argv[2]) : 1;
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  const int sms = prop.multiProcessorCount;
  const int table_mb = argc > 3 ? atoi(argv[3]) : 64;
  g_table_mb = table_mb;
  const size_t rows = (size_t)table_mb * 1024 * 1024 / row_bytes;
  const int n = wps > 4 ? 256 : 2048;
  std::vector<int> h((size_t)sms * wps * n);
  uint64_t s = 88172645463325252ull;
  for (auto& x : h) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; x = (int)(s % rows); }
  char* table; int* idx; float* out; long long* cyc;
  CK(cudaMalloc(&table, rows * row_bytes + 64)); CK(cudaMemset(table, 0, rows * row_bytes + 64));
  CK(cudaMalloc(&idx, h.size() * 4)); CK(cudaMemcpy(idx, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&out, 64)); CK(cudaMalloc(&cyc, sms * 8));
#define ALLU(M, NM) run<1, M>(table, idx, n, row_bytes, out, cyc, sms, wps, NM); run<2, M>(table, idx, n, row_bytes, out, cyc, sms, wps, NM); \
  run<4, M>(table, idx, n, row_bytes, out, cyc, sms, wps, NM); run<8, M>(table, idx, n, row_bytes, out, cyc, sms, wps, NM); run<16, M>(table, idx, n, row_bytes, out, cyc, sms, wps, NM);
  ALLU(0, "ldg")
  return 0;
}
