// Fused tail of the GCN forward / head of its backward: log-softmax over the class logits + NLL loss on the
// training vertices, and the gradient of that loss with respect to the logits.
// Replaces F.log_softmax(x, dim=1) (pygcn/gcn6.py:248), F.nll_loss(output[idx_train], labels[idx_train])
// (gcn6.py:427) and what autograd derives from them (index backward into a zero matrix, log-softmax backward,
// the slice/pad copies around the padded logits): ~4 GB of elementwise traffic per products-shaped epoch as
// separate ATen kernels, 1.4 GB here.
//   gcnb_logsoftmax_nll_fwd   logp[r, :] = logits[r, :] - logsumexp(logits[r, :]) for every vertex (the model's
//                             output, same as the reference's), loss = -mean_{i in idx_train} logp[i, label_i].
//                             The loss is summed by ONE block in a fixed order: bit-reproducible.
//   gcnb_logsoftmax_nll_bwd   dlogits[r, c] = g * (exp(logp[r, c]) - [c == label_r]) / n_train for training
//                             rows, 0 for all others and for the padding columns c >= ncls (the logits come
//                             out of the aggregation SpMM padded to a multiple of 4 columns; its transposed
//                             product takes this gradient as is).
// The class count is small (7 .. 47 .. a few hundred): 8 lanes per row, 4 rows per warp.
#include <cuda_runtime.h>
#include <stdint.h>

#include "common.h"

namespace gcnb {

// 8 lanes per row, 4 rows per warp (the class count is a few dozen: a whole warp per row leaves most lanes idle
// and serialises two shuffle reductions per 47 values — 0.70 ms on the products-shaped epoch, 0.93 GB)
constexpr int kLaneGroup = 8;

__global__ void __launch_bounds__(256)
logsoftmax_rows_kernel(const float* __restrict__ logits, int64_t ldl, int64_t n, int ncls,
                       float* __restrict__ logp, int64_t ldo) {
  const int lane = threadIdx.x & 31, lg = lane & (kLaneGroup - 1);
  const int64_t row = ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * (32 / kLaneGroup) + lane / kLaneGroup;
  const bool ok = row < n;
  const float* x = logits + (ok ? row : 0) * ldl;
  float mx = -INFINITY;
  if (ok)
    for (int c = lg; c < ncls; c += kLaneGroup) mx = fmaxf(mx, x[c]);
#pragma unroll
  for (int o = kLaneGroup / 2; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float s = 0.f;
  if (ok)
    for (int c = lg; c < ncls; c += kLaneGroup) s += expf(x[c] - mx);
#pragma unroll
  for (int o = kLaneGroup / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);   // xor butterfly: same value in every lane of the group
  if (!ok) return;
  const float lse = mx + logf(s);
  float* y = logp + row * ldo;
  for (int c = lg; c < ncls; c += kLaneGroup) y[c] = x[c] - lse;
}

// loss = -(1/denom) * sum_i logp[idx[i], labels[idx[i]]], one block, fixed summation order (denom = number of
// training vertices; on a row partition: the global count, each rank summing its own rows)
__global__ void __launch_bounds__(1024)
nll_mean_kernel(const float* __restrict__ logp, int64_t ldo, const int64_t* __restrict__ labels,
                const int64_t* __restrict__ idx, int64_t t, float inv_denom, float* __restrict__ loss) {
  __shared__ float part[1024];
  float s = 0.f;
  for (int64_t i = threadIdx.x; i < t; i += 1024) {
    const int64_t r = idx[i];
    s -= logp[r * ldo + labels[r]];
  }
  part[threadIdx.x] = s;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) part[threadIdx.x] += part[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) loss[0] = part[0] * inv_denom;
}

__global__ void __launch_bounds__(256)
logsoftmax_nll_bwd_kernel(const float* __restrict__ logp, int64_t ldo, const int64_t* __restrict__ labels,
                          const uint8_t* __restrict__ is_train, const float* __restrict__ gloss, float inv_t,
                          int64_t n, int ncls, int ncols, float* __restrict__ dlogits, int64_t ldd) {
  const int lane = threadIdx.x & 31, lg = lane & (kLaneGroup - 1);
  const int64_t row = ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * (32 / kLaneGroup) + lane / kLaneGroup;
  if (row >= n) return;
  float* d = dlogits + row * ldd;
  if (!is_train[row]) {
    for (int c = lg; c < ncols; c += kLaneGroup) d[c] = 0.f;
    return;
  }
  const float g = gloss[0] * inv_t;
  const int lab = (int)labels[row];
  const float* y = logp + row * ldo;
  for (int c = lg; c < ncols; c += kLaneGroup)
    d[c] = c < ncls ? g * (expf(y[c]) - (c == lab ? 1.f : 0.f)) : 0.f;
}

}  // namespace gcnb

using namespace gcnb;

extern "C" int gcnb_logsoftmax_nll_fwd(const float* logits, int64_t ldl, int64_t n, int ncls, const int64_t* labels,
                                       const int64_t* idx_train, int64_t n_train, int64_t denom, float* logp,
                                       int64_t ldo, float* loss, void* stream_) {
  GCNB_REQUIRE(n >= 0 && ncls >= 1 && n_train >= 0, "bad sizes (n=%lld ncls=%d n_train=%lld)", (long long)n, ncls,
               (long long)n_train);
  GCNB_REQUIRE(ldl >= ncls && ldo >= ncls, "leading dimension smaller than the class count");
  GCNB_REQUIRE(n == 0 || (logits && logp), "null logits / logp");
  GCNB_REQUIRE(loss != nullptr && (n_train == 0 || (labels && idx_train)), "null loss / labels / idx_train");
  int rc = require_gpu();
  if (rc != GCNB_OK) return rc;
  const int dev = pointer_device(logp ? (const void*)logp : (const void*)loss);
  GCNB_REQUIRE(dev >= 0, "logp is not a device pointer");
  DeviceGuard guard(dev);
  GCNB_CHECK_CUDA(guard.status);
  cudaStream_t stream = (cudaStream_t)stream_;
  if (n > 0) {
    logsoftmax_rows_kernel<<<(unsigned)((n + 31) / 32), 256, 0, stream>>>(logits, ldl, n, ncls, logp, ldo);
    GCNB_CHECK_CUDA(cudaGetLastError());
  }
  nll_mean_kernel<<<1, 1024, 0, stream>>>(logp, ldo, labels, idx_train, n_train, denom > 0 ? 1.f / (float)denom : 0.f, loss);
  GCNB_CHECK_CUDA(cudaGetLastError());
  return GCNB_OK;
}

extern "C" int gcnb_logsoftmax_nll_bwd(const float* logp, int64_t ldo, const int64_t* labels, const uint8_t* is_train,
                                       const float* gloss, int64_t denom, int64_t n, int ncls, int ncols,
                                       float* dlogits, int64_t ldd, void* stream_) {
  GCNB_REQUIRE(n >= 0 && ncls >= 1 && ncols >= ncls && ldd >= ncols && ldo >= ncls, "bad sizes");
  if (n == 0) return GCNB_OK;
  GCNB_REQUIRE(logp && labels && is_train && gloss && dlogits, "null pointer");
  int rc = require_gpu();
  if (rc != GCNB_OK) return rc;
  const int dev = pointer_device(dlogits);
  GCNB_REQUIRE(dev >= 0, "dlogits is not a device pointer");
  DeviceGuard guard(dev);
  GCNB_CHECK_CUDA(guard.status);
  logsoftmax_nll_bwd_kernel<<<(unsigned)((n + 31) / 32), 256, 0, (cudaStream_t)stream_>>>(
      logp, ldo, labels, is_train, gloss, denom > 0 ? 1.f / (float)denom : 0.f, n, ncls, ncols, dlogits, ldd);
  GCNB_CHECK_CUDA(cudaGetLastError());
  return GCNB_OK;
}
