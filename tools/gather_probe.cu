// Microbenchmark: what row-gather rate can the SMs sustain out of L2 / HBM on this B200?
// Sets the physical target for the SpMM rows kernels (every non-zero is one gathered feature row).
//   gather_probe <table_MB> <row_bytes> <warps_per_cta> <ctas_per_sm> <unroll> [window_rows]
// Each warp reads rows table[idx[i]] (lanes own 16-byte slices) for a private random index stream.
// window_rows > 0: indices are drawn within +-window of a position that sweeps through the table (models a
// locality-ordered graph); 0: uniform over the whole table.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

template <int U, int VB>   // VB = bytes per lane per load (16, 8, 4)
__global__ void gather_kernel(const char* __restrict__ table, const int* __restrict__ idx, int per_warp,
                              int row_bytes, int act_bytes, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int* my = idx + (size_t)gw * per_warp;
  const bool act = lane * VB < act_bytes;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int i = 0; i < per_warp; i += 32) {
    const int mine = my[i + lane];
    for (int j = 0; j < 32; j += U) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int r = __shfl_sync(0xffffffffu, mine, j + u);
        const char* p = table + (size_t)r * row_bytes + lane * VB;
        v[u] = make_float4(0, 0, 0, 0);
        if (act) {
          if (VB == 16) v[u] = __ldg(reinterpret_cast<const float4*>(p));
          else if (VB == 8) { float2 t = __ldg(reinterpret_cast<const float2*>(p)); v[u].x = t.x; v[u].y = t.y; }
          else v[u].x = __ldg(reinterpret_cast<const float*>(p));
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) { acc[0] += v[u].x; acc[1] += v[u].y; acc[2] += v[u].z; acc[3] += v[u].w; }
    }
  }
  if (acc[0] + acc[1] + acc[2] + acc[3] == 12345.678f) out[gw] = acc[0];
}

// plain streaming read of the table (L2 / HBM sequential bandwidth reference)
__global__ void stream_kernel(const float4* __restrict__ t, size_t n4, float* out) {
  float4 a = make_float4(0, 0, 0, 0);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    float4 v = __ldg(t + i); a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
  }
  if (a.x + a.y + a.z + a.w == 12345.678f) out[0] = a.x;
}

template <int U, int VB>
float run(const char* table, const int* idx, int per_warp, int row_bytes, int act_bytes, float* out, int grid, int block) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 2; ++i) gather_kernel<U, VB><<<grid, block>>>(table, idx, per_warp, row_bytes, act_bytes, out);
  CK(cudaDeviceSynchronize());
  cudaEventRecord(a);
  for (int i = 0; i < 5; ++i) gather_kernel<U, VB><<<grid, block>>>(table, idx, per_warp, row_bytes, act_bytes, out);
  cudaEventRecord(b); CK(cudaDeviceSynchronize());
  float ms; cudaEventElapsedTime(&ms, a, b); return ms / 5;
}

int main(int argc, char** argv) {
  const size_t table_mb = argc > 1 ? atoi(argv[1]) : 64;
  const int row_bytes = argc > 2 ? atoi(argv[2]) : 400;
  const int wpc = argc > 3 ? atoi(argv[3]) : 8;
  const int cps = argc > 4 ? atoi(argv[4]) : 4;
  const int unroll = argc > 5 ? atoi(argv[5]) : 8;
  const long window = argc > 6 ? atol(argv[6]) : 0;
  int vb = argc > 7 ? atoi(argv[7]) : 16;
  const int act_bytes = argc > 8 ? atoi(argv[8]) : row_bytes;   // bytes of each row actually read (pitch = row_bytes)
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  const int sms = prop.multiProcessorCount;
  const size_t rows = table_mb * 1024 * 1024 / row_bytes;
  const int grid = sms * cps, block = wpc * 32;
  const int warps = grid * wpc;
  const int per_warp = 4096;
  std::vector<int> h((size_t)warps * per_warp);
  uint64_t s = 88172645463325252ull;
  auto rnd = [&]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; };
  for (int w = 0; w < warps; ++w)
    for (int i = 0; i < per_warp; ++i) {
      long r;
      if (window > 0) {
        // warp w sweeps its own contiguous slice of the table (like a contiguous row block of the matrix)
        const long centre = (long)((double)w / warps * rows) + (long)((double)i / per_warp * rows / warps);
        r = centre + (long)(rnd() % (2 * window + 1)) - window;
        r = ((r % (long)rows) + rows) % rows;
      } else r = rnd() % rows;
      h[(size_t)w * per_warp + i] = (int)r;
    }
  char* table; int* idx; float* out;
  CK(cudaMalloc(&table, rows * row_bytes + 64)); CK(cudaMemset(table, 0, rows * row_bytes + 64));
  CK(cudaMalloc(&idx, h.size() * 4)); CK(cudaMemcpy(idx, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&out, warps * 4 + 4));
  float ms = 0;
#define RUN(UU) (vb == 16 ? run<UU, 16>(table, idx, per_warp, row_bytes, act_bytes, out, grid, block) : vb == 8 ? run<UU, 8>(table, idx, per_warp, row_bytes, act_bytes, out, grid, block) : run<UU, 4>(table, idx, per_warp, row_bytes, act_bytes, out, grid, block))
  switch (unroll) { case 1: ms = RUN(1); break; case 2: ms = RUN(2); break; case 4: ms = RUN(4); break; case 8: ms = RUN(8); break; case 16: ms = RUN(16); break; default: ms = RUN(32); }
  const double gathered = (double)warps * per_warp;
  const double bytes = gathered * (act_bytes < 32 * vb ? act_bytes : 32 * vb);
  const double clk = prop.clockRate * 1e3;   // Hz (max SM clock)
  // streaming reference over the same table
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  const size_t n4 = rows * row_bytes / 16;
  stream_kernel<<<sms * 8, 512>>>((const float4*)table, n4, out); CK(cudaDeviceSynchronize());
  cudaEventRecord(a);
  for (int i = 0; i < 5; ++i) stream_kernel<<<sms * 8, 512>>>((const float4*)table, n4, out);
  cudaEventRecord(b); CK(cudaDeviceSynchronize());
  float sms_; cudaEventElapsedTime(&sms_, a, b); sms_ /= 5;
  printf("{\"table_mb\": %zu, \"row_bytes\": %d, \"warps_per_cta\": %d, \"ctas_per_sm\": %d, \"unroll\": %d, \"window\": %ld, \"vb\": %d, \"act_bytes\": %d, "
         "\"ms\": %.4f, \"gather_GBps\": %.1f, \"rows_per_us\": %.1f, \"cyc_per_row_per_sm\": %.2f, \"stream_GBps\": %.1f}\n",
         table_mb, row_bytes, wpc, cps, unroll, window, vb, act_bytes, ms, bytes / ms / 1e6, gathered / ms / 1e3,
         ms * 1e-3 * clk * sms / gathered, (double)n4 * 16 / sms_ / 1e6);
  return 0;
}
