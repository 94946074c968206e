// Device side of the SpMM engine: C[m x k] = A[m x n] (CSR, fp32 values) * B[n x k].
//
// Replaces the reference's flexspmm_cuda_* kernels (flexspmm.cu:17-498).  Not a port: the
// reference walks 8-row tile-segments handed out by per-SM atomic counters and atomically adds
// partial rows; here the matrix stays in (interleaved) CSR, every output element is produced by
// exactly one warp or one CTA, and nothing is atomic.
//
//   stream class (spmm_stream_kernel)        the default for feature rows wider than 16 vectors: a warp owns
//                                    a segment of ~768 consecutive non-zeros (whole short rows or a
//                                    piece of a long row) and streams it through a register ring of
//                                    in-flight feature-row gathers; (col,val) pairs arrive through a
//                                    per-warp shared-memory slab.
//   packed stream (spmm_stream_packed_kernel)  narrow feature rows on matrices with >= 24 nnz/row:
//                                    the same, with S = 2/4/8 non-zeros fetched per warp-wide load.
//   rows class  (spmm_rows_kernel)   narrow feature rows on short-row matrices: one lane-group of G
//                                    lanes per row, 32/G rows per warp.  (col,val) pairs are fetched G
//                                    at a time as one coalesced 8-byte-per-lane load and broadcast with
//                                    __shfl_sync; U independent gathers are in flight per lane.
//   block class (spmm_block_kernel)  companion of the rows class for rows with >= kLongRow non-zeros,
//                                    cut into items of at most kChunk non-zeros.  One CTA per item: the
//                                    item's (col,val) tile is staged into shared memory with one TMA
//                                    bulk copy (cp.async.bulk + mbarrier), the 8 warps stride over the
//                                    tile, partial rows are combined with warp shuffles and a
//                                    shared-memory pass in fixed order.
//   Rows cut into several pieces go through a partials buffer and spmm_reduce_kernel — still no
//   atomics, so results are bit-reproducible.  Dispatch: spmm.cu launch_vec; measurements behind every
//   choice: DESIGN.md §4.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace gcnb {

constexpr int kLongRow = 512;    // >= this many non-zeros: block class
constexpr int kChunk = 2048;     // non-zeros per block-class work item (16 KB of smem)
constexpr int kRowWarps = 4;     // warps per CTA, rows class
constexpr int kBlockWarps = 8;   // warps per CTA, block class

struct WorkItem {
  int32_t row;
  int32_t beg;   // first non-zero (index into cv)
  int32_t end;   // one past the last
  int32_t slot;  // -1: the item is the whole row, store to C; else index into partials
};

// ---- loads -----------------------------------------------------------------------------------
// A stream (col,val): read once per column tile -> keep it out of L1.
__device__ __forceinline__ int2 ld_stream_int2(const int2* p) {
  int2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.s32 {%0, %1}, [%2];"
               : "=r"(r.x), "=r"(r.y)
               : "l"(p));
  return r;
}

// L2 eviction priorities (createpolicy + .L2::cache_hint; the policy rides in the memory descriptor, so a
// hinted access costs no extra instruction).  The (col,val) stream and the C rows are touched once:
// evict_first keeps them from pushing the re-used feature rows out of the 126 MB L2.
__device__ __forceinline__ uint64_t l2_policy(int kind) {   // 0 normal, 1 evict_first, 2 evict_last
  uint64_t pol;
  if (kind == 1) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  else if (kind == 2) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  else asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ int2 ld_stream_int2(const int2* p, uint64_t pol) {
  int2 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.s32 {%0, %1}, [%2], %3;"
               : "=r"(r.x), "=r"(r.y)
               : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ void ldg_hint(const void* p, uint64_t pol, uint4& v) {
  asm volatile("ld.global.nc.L2::cache_hint.v4.b32 {%0, %1, %2, %3}, [%4], %5;"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p), "l"(pol));
}
// 256-bit load (sm_100: LDG.E.ENL2.256): one lane brings 32 bytes, so a 400-byte feature row needs 13 lanes
// and ONE warp-wide request can carry two rows.  The address must be 32-byte aligned.
__device__ __forceinline__ void ldg_hint256(const void* p, uint64_t pol, float (&w)[8]) {
  asm volatile("ld.global.nc.L2::cache_hint.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8], %9;"
               : "=f"(w[0]), "=f"(w[1]), "=f"(w[2]), "=f"(w[3]), "=f"(w[4]), "=f"(w[5]), "=f"(w[6]), "=f"(w[7])
               : "l"(p), "l"(pol));
}
__device__ __forceinline__ void ldg256(const void* p, float (&w)[8]) {
  asm volatile("ld.global.nc.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=f"(w[0]), "=f"(w[1]), "=f"(w[2]), "=f"(w[3]), "=f"(w[4]), "=f"(w[5]), "=f"(w[6]), "=f"(w[7])
               : "l"(p));
}
__device__ __forceinline__ void ldg_hint(const void* p, uint64_t pol, uint2& v) {
  asm volatile("ld.global.nc.L2::cache_hint.v2.b32 {%0, %1}, [%2], %3;" : "=r"(v.x), "=r"(v.y) : "l"(p), "l"(pol));
}
__device__ __forceinline__ void ldg_hint(const void* p, uint64_t pol, uint32_t& v) {
  asm volatile("ld.global.nc.L2::cache_hint.b32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
}
__device__ __forceinline__ void ldg_hint(const void* p, uint64_t pol, unsigned short& v) {
  asm volatile("ld.global.nc.L2::cache_hint.u16 %0, [%1], %2;" : "=h"(v) : "l"(p), "l"(pol));
}

template <typename TB, int VEC>
struct VecLoad;

template <>
struct VecLoad<float, 4> {
  static __device__ __forceinline__ void load(const float* p, float (&o)[4]) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(p));
    o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w;
  }
  // coherent load (C in accumulate mode was written by the previous kernel, not read-only data)
  static __device__ __forceinline__ void load_plain(const float* p, float (&o)[4]) {
    const float4 v = *reinterpret_cast<const float4*>(p);
    o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w;
  }
};
template <>
struct VecLoad<float, 8> {
  static __device__ __forceinline__ void load(const float* p, float (&o)[8]) { ldg256(p, o); }
  static __device__ __forceinline__ void load_plain(const float* p, float (&o)[8]) {
    const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
  }
};
template <>
struct VecLoad<float, 2> {
  static __device__ __forceinline__ void load(const float* p, float (&o)[2]) {
    const float2 v = __ldg(reinterpret_cast<const float2*>(p));
    o[0] = v.x; o[1] = v.y;
  }
  static __device__ __forceinline__ void load_plain(const float* p, float (&o)[2]) {
    const float2 v = *reinterpret_cast<const float2*>(p);
    o[0] = v.x; o[1] = v.y;
  }
};
template <>
struct VecLoad<float, 1> {
  static __device__ __forceinline__ void load(const float* p, float (&o)[1]) { o[0] = __ldg(p); }
  static __device__ __forceinline__ void load_plain(const float* p, float (&o)[1]) { o[0] = *p; }
};

__device__ __forceinline__ void bf16x2_to_f32(uint32_t w, float& lo, float& hi) {
  lo = __uint_as_float(w << 16);
  hi = __uint_as_float(w & 0xffff0000u);
}
template <>
struct VecLoad<__nv_bfloat16, 8> {
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&o)[8]) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
    bf16x2_to_f32(v.x, o[0], o[1]); bf16x2_to_f32(v.y, o[2], o[3]);
    bf16x2_to_f32(v.z, o[4], o[5]); bf16x2_to_f32(v.w, o[6], o[7]);
  }
};
template <>
struct VecLoad<__nv_bfloat16, 4> {
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&o)[4]) {
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
    bf16x2_to_f32(v.x, o[0], o[1]); bf16x2_to_f32(v.y, o[2], o[3]);
  }
};
template <>
struct VecLoad<__nv_bfloat16, 2> {
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&o)[2]) {
    const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(p));
    bf16x2_to_f32(v, o[0], o[1]);
  }
};
template <>
struct VecLoad<__nv_bfloat16, 1> {
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&o)[1]) {
    const unsigned short v = __ldg(reinterpret_cast<const unsigned short*>(p));
    o[0] = __uint_as_float(((uint32_t)v) << 16);
  }
};

// Raw (unconverted) vector loads for kernels that keep gathered rows in flight in registers: the
// bf16 -> fp32 conversion must not sit next to the load (it would wait for the data at once and
// serialise the gathers); it happens when the slot is consumed.
template <typename TB, int VEC>
struct RawVec;
template <int VEC>
struct RawVec<float, VEC> {
  float w[VEC];
  __device__ __forceinline__ void load(const float* p) { VecLoad<float, VEC>::load(p, w); }
  __device__ __forceinline__ void load(const float* p, uint64_t pol) {
    if constexpr (VEC == 8) {
      ldg_hint256(p, pol, w);
    } else if constexpr (VEC == 4) {
      uint4 v; ldg_hint(p, pol, v);
      w[0] = __uint_as_float(v.x); w[1] = __uint_as_float(v.y); w[2] = __uint_as_float(v.z); w[3] = __uint_as_float(v.w);
    } else if constexpr (VEC == 2) {
      uint2 v; ldg_hint(p, pol, v);
      w[0] = __uint_as_float(v.x); w[1] = __uint_as_float(v.y);
    } else {
      uint32_t v; ldg_hint(p, pol, v);
      w[0] = __uint_as_float(v);
    }
  }
  __device__ __forceinline__ void zero() {
#pragma unroll
    for (int x = 0; x < VEC; ++x) w[x] = 0.f;
  }
  __device__ __forceinline__ void get(float (&o)[VEC]) const {
#pragma unroll
    for (int x = 0; x < VEC; ++x) o[x] = w[x];
  }
};
template <>
struct RawVec<__nv_bfloat16, 8> {
  uint4 w;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { w = __ldg(reinterpret_cast<const uint4*>(p)); }
  __device__ __forceinline__ void load(const __nv_bfloat16* p, uint64_t pol) { ldg_hint(p, pol, w); }
  __device__ __forceinline__ void zero() { w = make_uint4(0, 0, 0, 0); }
  __device__ __forceinline__ void get(float (&o)[8]) const {
    bf16x2_to_f32(w.x, o[0], o[1]); bf16x2_to_f32(w.y, o[2], o[3]);
    bf16x2_to_f32(w.z, o[4], o[5]); bf16x2_to_f32(w.w, o[6], o[7]);
  }
};
template <>
struct RawVec<__nv_bfloat16, 4> {
  uint2 w;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { w = __ldg(reinterpret_cast<const uint2*>(p)); }
  __device__ __forceinline__ void load(const __nv_bfloat16* p, uint64_t pol) { ldg_hint(p, pol, w); }
  __device__ __forceinline__ void zero() { w = make_uint2(0, 0); }
  __device__ __forceinline__ void get(float (&o)[4]) const {
    bf16x2_to_f32(w.x, o[0], o[1]); bf16x2_to_f32(w.y, o[2], o[3]);
  }
};
template <>
struct RawVec<__nv_bfloat16, 2> {
  uint32_t w;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { w = __ldg(reinterpret_cast<const uint32_t*>(p)); }
  __device__ __forceinline__ void load(const __nv_bfloat16* p, uint64_t pol) { ldg_hint(p, pol, w); }
  __device__ __forceinline__ void zero() { w = 0; }
  __device__ __forceinline__ void get(float (&o)[2]) const { bf16x2_to_f32(w, o[0], o[1]); }
};
template <>
struct RawVec<__nv_bfloat16, 1> {
  unsigned short w;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { w = __ldg(reinterpret_cast<const unsigned short*>(p)); }
  __device__ __forceinline__ void load(const __nv_bfloat16* p, uint64_t pol) { ldg_hint(p, pol, w); }
  __device__ __forceinline__ void zero() { w = 0; }
  __device__ __forceinline__ void get(float (&o)[1]) const { o[0] = __uint_as_float(((uint32_t)w) << 16); }
};

// C is written once and not re-read by this kernel: streaming stores.
template <int VEC>
__device__ __forceinline__ void store_vec(float* p, const float (&v)[VEC]) {
  if constexpr (VEC % 4 == 0) {
#pragma unroll
    for (int i = 0; i < VEC; i += 4)
      __stcs(reinterpret_cast<float4*>(p + i), make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]));
  } else if constexpr (VEC == 2) {
    __stcs(reinterpret_cast<float2*>(p), make_float2(v[0], v[1]));
  } else {
    __stcs(p, v[0]);
  }
}

template <int VEC>
__device__ __forceinline__ void store_vec(float* p, const float (&v)[VEC], uint64_t pol) {
  if constexpr (VEC % 4 == 0) {
#pragma unroll
    for (int i = 0; i < VEC; i += 4)
      asm volatile("st.global.L1::no_allocate.L2::cache_hint.v4.f32 [%0], {%1, %2, %3, %4}, %5;"
                   ::"l"(p + i), "f"(v[i]), "f"(v[i + 1]), "f"(v[i + 2]), "f"(v[i + 3]), "l"(pol) : "memory");
  } else if constexpr (VEC == 2) {
    asm volatile("st.global.L1::no_allocate.L2::cache_hint.v2.f32 [%0], {%1, %2}, %3;"
                 ::"l"(p), "f"(v[0]), "f"(v[1]), "l"(pol) : "memory");
  } else {
    asm volatile("st.global.L1::no_allocate.L2::cache_hint.f32 [%0], %1, %2;" ::"l"(p), "f"(v[0]), "l"(pol) : "memory");
  }
}

// existing C values in accumulate mode: coherent loads (written by the previous kernel), same vector
// widths as store_vec
template <int VEC>
__device__ __forceinline__ void load_c(const float* p, float (&v)[VEC]) {
  if constexpr (VEC % 4 == 0) {
#pragma unroll
    for (int i = 0; i < VEC; i += 4) {
      const float4 q = *reinterpret_cast<const float4*>(p + i);
      v[i] = q.x; v[i + 1] = q.y; v[i + 2] = q.z; v[i + 3] = q.w;
    }
  } else if constexpr (VEC == 2) {
    const float2 q = *reinterpret_cast<const float2*>(p);
    v[0] = q.x; v[1] = q.y;
  } else {
    v[0] = *p;
  }
}

// number of the VEC columns starting at `coff` that exist (k is a multiple of 4 whenever VEC is 8: a lane of the
// 256-bit kernels may own 4 real columns and 4 of the row's padding)
template <int VEC>
__device__ __forceinline__ int cols_valid(int coff, int k) {
  if constexpr (VEC == 8) return coff + 8 > k ? 4 : 8;
  else return VEC;
}
template <int VEC>
__device__ __forceinline__ void store_row_vec(float* p, const float (&v)[VEC], int nval, uint64_t pol) {
  if constexpr (VEC == 8) {
    const float lo[4] = {v[0], v[1], v[2], v[3]}, hi[4] = {v[4], v[5], v[6], v[7]};
    store_vec<4>(p, lo, pol);
    if (nval == 8) store_vec<4>(p + 4, hi, pol);
  } else {
    store_vec<VEC>(p, v, pol);
  }
}
template <int VEC>
__device__ __forceinline__ void store_row_vec(float* p, const float (&v)[VEC], int nval) {
  if constexpr (VEC == 8) {
    const float lo[4] = {v[0], v[1], v[2], v[3]}, hi[4] = {v[4], v[5], v[6], v[7]};
    store_vec<4>(p, lo);
    if (nval == 8) store_vec<4>(p + 4, hi);
  } else {
    store_vec<VEC>(p, v);
  }
}

__device__ __forceinline__ float epilogue(float x, const float* bias, int col, int epi) {
  if (epi & GCNB_EPI_BIAS) x += __ldg(bias + col);
  if (epi & GCNB_EPI_RELU) x = fmaxf(x, 0.f);
  return x;
}

template <int NV, int VEC>
__host__ __device__ constexpr int unroll_for() {
  return (NV * VEC >= 16) ? 2 : (NV * VEC >= 8 ? 4 : 8);
}

// ---- rows class --------------------------------------------------------------------------------
template <typename TB, int VEC, int G, int NV>
__global__ void __launch_bounds__(kRowWarps * 32)
spmm_rows_kernel(const int32_t* __restrict__ rowptr, const int2* __restrict__ cv,
                 const TB* __restrict__ B, int64_t ldb, float* __restrict__ C, int64_t ldc,
                 int64_t m, int k, const float* __restrict__ bias, int epi, int long_thresh) {
  constexpr int RPW = 32 / G;
  constexpr int U0 = unroll_for<NV, VEC>();
  constexpr int U = U0 < G ? U0 : G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int lig = lane & (G - 1), grp = lane / G;
  const int64_t row = ((int64_t)blockIdx.x * kRowWarps + warp) * RPW + grp;
  const int colbase = blockIdx.y * (G * NV * VEC);

  int beg = 0, len = 0;
  bool own = false;
  if (row < m) {
    beg = __ldg(rowptr + row);
    len = __ldg(rowptr + row + 1) - beg;
    own = len < long_thresh;
    if (!own) len = 0;  // the block class computes this row
  }
  int coff[NV];
  bool cok[NV];
#pragma unroll
  for (int t = 0; t < NV; ++t) {
    coff[t] = colbase + (lig + t * G) * VEC;
    cok[t] = coff[t] < k;
  }
  float acc[NV][VEC];
#pragma unroll
  for (int t = 0; t < NV; ++t)
#pragma unroll
    for (int x = 0; x < VEC; ++x) acc[t][x] = 0.f;

  int maxlen = len;  // warp-uniform trip count keeps the lane groups converged
  if constexpr (G < 32) {
#pragma unroll
    for (int o = G; o < 32; o <<= 1) maxlen = max(maxlen, __shfl_xor_sync(0xffffffffu, maxlen, o));
  }

  for (int base = 0; base < maxlen; base += G) {
    int2 e = make_int2(0, 0);
    if (base + lig < len) e = ld_stream_int2(cv + (int64_t)beg + base + lig);
    const int cnt = min(len - base, G);       // group-uniform, may be <= 0
    const int wcnt = min(maxlen - base, G);   // warp-uniform
#pragma unroll
    for (int j0 = 0; j0 < G; j0 += U) {
      if (j0 >= wcnt) break;
      float bv[U][NV][VEC];
      float av[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int c = __shfl_sync(0xffffffffu, e.x, j0 + u, G);
        const float v = __int_as_float(__shfl_sync(0xffffffffu, e.y, j0 + u, G));
        const bool p = (j0 + u) < cnt;
        av[u] = p ? v : 0.f;
        const TB* bp = B + (int64_t)c * ldb;
#pragma unroll
        for (int t = 0; t < NV; ++t) {
          if (p && cok[t]) {
            VecLoad<TB, VEC>::load(bp + coff[t], bv[u][t]);
          } else {
#pragma unroll
            for (int x = 0; x < VEC; ++x) bv[u][t][x] = 0.f;
          }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
#pragma unroll
        for (int t = 0; t < NV; ++t)
#pragma unroll
          for (int x = 0; x < VEC; ++x) acc[t][x] = fmaf(av[u], bv[u][t][x], acc[t][x]);
    }
  }

  if (own) {
    float* crow = C + row * ldc;
#pragma unroll
    for (int t = 0; t < NV; ++t) {
      if (cok[t]) {
        float o[VEC];
        if (epi & GCNB_EPI_ACCUM) {
#pragma unroll
          for (int x = 0; x < VEC; ++x) acc[t][x] += crow[coff[t] + x];
        }
#pragma unroll
        for (int x = 0; x < VEC; ++x) o[x] = epilogue(acc[t][x], bias, coff[t] + x, epi);
        store_vec<VEC>(crow + coff[t], o);
      }
    }
  }
}

// ---- rows class, wide feature rows: one warp per row -------------------------------------------
// Used when a feature row needs more than 16 lanes' worth of vectors.  Differences from the
// lane-group kernel above, all aimed at the instruction count per non-zero (the first version spent
// 32 issue slots per non-zero, two of them shuffles that share the LSU data pipe with the gather):
//   * the warp's (col,val) pairs are staged 32 at a time in a private shared-memory slab and read
//     back two pairs per LDS.128 broadcast — 0.5 instructions per non-zero instead of 2 shuffles;
//   * the gather address is one IMAD.WIDE: per-lane base pointer + col * (row pitch in bytes);
//   * full groups of U non-zeros run without predicates, the tail is peeled;
//   * multiply-adds are issued as packed fma.rn.f32x2 (two lanes of fp32 per instruction).
__device__ __forceinline__ void ffma2(float& c0, float& c1, float a, float b0, float b1) {
  unsigned long long c, aa, bb;
  asm("mov.b64 %0, {%1, %2};" : "=l"(c) : "f"(c0), "f"(c1));
  asm("mov.b64 %0, {%1, %1};" : "=l"(aa) : "f"(a));
  asm("mov.b64 %0, {%1, %2};" : "=l"(bb) : "f"(b0), "f"(b1));
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(c) : "l"(aa), "l"(bb));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(c0), "=f"(c1) : "l"(c));
}

template <int NV, int VEC>
__device__ __forceinline__ void fma_row(float (&acc)[NV][VEC], float a, const float (&b)[NV][VEC]) {
#pragma unroll
  for (int t = 0; t < NV; ++t) {
    if constexpr (VEC % 2 == 0) {
#pragma unroll
      for (int x = 0; x < VEC; x += 2) ffma2(acc[t][x], acc[t][x + 1], a, b[t][x], b[t][x + 1]);
    } else {
#pragma unroll
      for (int x = 0; x < VEC; ++x) acc[t][x] = fmaf(a, b[t][x], acc[t][x]);
    }
  }
}

constexpr int kWarpRowWarps = 8;

template <typename TB, int VEC, int NV, bool PIPE>
__global__ void __launch_bounds__(kWarpRowWarps * 32)
spmm_warp_row_kernel(const int32_t* __restrict__ rowptr, const int2* __restrict__ cv,
                     const TB* __restrict__ B, uint32_t ldb_bytes, float* __restrict__ C, int64_t ldc,
                     int64_t m, int k, const float* __restrict__ bias, int epi, int long_thresh) {
  constexpr int U0 = (NV * VEC >= 16) ? 2 : (NV * VEC >= 8 ? 4 : 8);
  constexpr int U = (PIPE && U0 > 2) ? U0 / 2 : U0;   // two register buffers when pipelined
  __shared__ alignas(16) int2 s_cv[kWarpRowWarps][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t row = (int64_t)blockIdx.x * kWarpRowWarps + warp;
  if (row >= m) return;   // warp-uniform; no block-wide barrier below
  const int colbase = blockIdx.y * (32 * NV * VEC);
  const int beg = __ldg(rowptr + row);
  int len = __ldg(rowptr + row + 1) - beg;
  if (len >= long_thresh) return;   // the block class computes this row
  if (len == 0 && epi == GCNB_EPI_ACCUM) return;   // C += 0: nothing to do

  bool cok[NV];
  int coff[NV];
#pragma unroll
  for (int t = 0; t < NV; ++t) {
    coff[t] = colbase + (lane + t * 32) * VEC;
    cok[t] = coff[t] < k;
  }
  // per-lane base: out-of-range lanes are parked on column 0 and never load
  const char* bl = reinterpret_cast<const char*>(B + (cok[0] ? coff[0] : 0));
  // accumulate mode starts from the existing C row: that load depends on nothing but the row id, so
  // it is issued first and arrives under the rowptr -> (col,val) -> feature-row chain
  float acc[NV][VEC];
#pragma unroll
  for (int t = 0; t < NV; ++t) {
    if ((epi & GCNB_EPI_ACCUM) && cok[t]) {
      load_c<VEC>(C + row * ldc + coff[t], acc[t]);
    } else {
#pragma unroll
      for (int x = 0; x < VEC; ++x) acc[t][x] = 0.f;
    }
  }

  int2* slab = s_cv[warp];
  const int2* src = cv + beg;
  int2 e = make_int2(0, 0);
  if (lane < len) e = ld_stream_int2(src + lane);
  for (int base = 0; base < len; base += 32) {
    __syncwarp();               // everyone is done reading the previous slab
    slab[lane] = e;
    __syncwarp();
    if (base + 32 + lane < len) e = ld_stream_int2(src + base + 32 + lane);   // prefetch the next slab
    const int cnt = min(32, len - base);
    int j = 0;
    if constexpr (PIPE) {
      // software pipeline: the gathers of group g+1 are in flight while group g is multiplied
      // (ptxas otherwise sinks every load next to its first use and leaves ~2 loads in flight)
      float bvA[U][NV][VEC], bvB[U][NV][VEC];
      float avA[U], avB[U];
      auto gather = [&](int j0, float (&bv)[U][NV][VEC], float (&av)[U]) {
#pragma unroll
        for (int u = 0; u < U; u += 2) {
          const int4 q = *reinterpret_cast<const int4*>(slab + j0 + u);
          av[u] = __int_as_float(q.y);
          av[u + 1] = __int_as_float(q.w);
          const char* p0 = bl + (uint64_t)(uint32_t)q.x * ldb_bytes;
          const char* p1 = bl + (uint64_t)(uint32_t)q.z * ldb_bytes;
#pragma unroll
          for (int t = 0; t < NV; ++t) {
            if (cok[t]) {
              VecLoad<TB, VEC>::load(reinterpret_cast<const TB*>(p0) + t * 32 * VEC, bv[u][t]);
              VecLoad<TB, VEC>::load(reinterpret_cast<const TB*>(p1) + t * 32 * VEC, bv[u + 1][t]);
            } else {
#pragma unroll
              for (int x = 0; x < VEC; ++x) { bv[u][t][x] = 0.f; bv[u + 1][t][x] = 0.f; }
            }
          }
        }
      };
      const int groups = cnt / U;
      if (groups > 0) gather(0, bvA, avA);
      int g = 0;
      for (; g + 2 <= groups; g += 2) {
        gather((g + 1) * U, bvB, avB);
#pragma unroll
        for (int u = 0; u < U; ++u) fma_row<NV, VEC>(acc, avA[u], bvA[u]);
        if (g + 2 < groups) gather((g + 2) * U, bvA, avA);
#pragma unroll
        for (int u = 0; u < U; ++u) fma_row<NV, VEC>(acc, avB[u], bvB[u]);
      }
      if (g < groups) {
#pragma unroll
        for (int u = 0; u < U; ++u) fma_row<NV, VEC>(acc, avA[u], bvA[u]);
      }
      j = groups * U;
    } else {
    for (; j + U <= cnt; j += U) {
      float bv[U][NV][VEC];
      float av[U];
#pragma unroll
      for (int u = 0; u < U; u += 2) {
        const int4 q = *reinterpret_cast<const int4*>(slab + j + u);   // two pairs, broadcast
        av[u] = __int_as_float(q.y);
        av[u + 1] = __int_as_float(q.w);
        const char* p0 = bl + (uint64_t)(uint32_t)q.x * ldb_bytes;
        const char* p1 = bl + (uint64_t)(uint32_t)q.z * ldb_bytes;
#pragma unroll
        for (int t = 0; t < NV; ++t) {
          if (cok[t]) {
            VecLoad<TB, VEC>::load(reinterpret_cast<const TB*>(p0) + t * 32 * VEC, bv[u][t]);
            VecLoad<TB, VEC>::load(reinterpret_cast<const TB*>(p1) + t * 32 * VEC, bv[u + 1][t]);
          } else {
#pragma unroll
            for (int x = 0; x < VEC; ++x) { bv[u][t][x] = 0.f; bv[u + 1][t][x] = 0.f; }
          }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) fma_row<NV, VEC>(acc, av[u], bv[u]);
    }
    }
    for (; j < cnt; ++j) {   // peeled tail, one non-zero at a time
      const int2 q = slab[j];
      const char* p0 = bl + (uint64_t)(uint32_t)q.x * ldb_bytes;
      float bv[NV][VEC];
#pragma unroll
      for (int t = 0; t < NV; ++t) {
        if (cok[t]) {
          VecLoad<TB, VEC>::load(reinterpret_cast<const TB*>(p0) + t * 32 * VEC, bv[t]);
        } else {
#pragma unroll
          for (int x = 0; x < VEC; ++x) bv[t][x] = 0.f;
        }
      }
      fma_row<NV, VEC>(acc, __int_as_float(q.y), bv);
    }
  }

  float* crow = C + row * ldc;
#pragma unroll
  for (int t = 0; t < NV; ++t) {
    if (cok[t]) {
      float o[VEC];
#pragma unroll
      for (int x = 0; x < VEC; ++x) o[x] = epilogue(acc[t][x], bias, coff[t] + x, epi);
      store_vec<VEC>(crow + coff[t], o);
    }
  }
}

// ---- rows class, wide feature rows: non-zero stream over row segments ---------------------------
// Measured on the B200 (tools/gather_probe, tools/mlp_probe; profiles/r01_b_*.jsonl):
//   * an SM pulls one gathered 400-byte feature row out of L2 every ~6.1 cycles at best (its
//     64 B/cycle L2 port), but only with >= 32 warps resident;
//   * ONE warp completes a gathered row every ~97 cycles (L2 hit) / ~295 cycles (HBM) no matter how
//     many independent loads it has issued (U = 1..16 gave the same rate), so the parallelism has
//     to come from the number of resident warps, not from deep per-warp unrolling;
//   * the warp-per-row kernel above spends ~370 cycles per non-zero per warp: three dependent
//     latencies per matrix row (rowptr -> (col,val) -> feature rows), a drain at every row end and
//     ragged row lengths inside a CTA.
// Here a warp owns a SEGMENT of consecutive rows (~kSegNnz non-zeros, cut on the host at plan time)
// and treats their (col,val) pairs as one continuous stream:
//   * a ring of U register slots holds the gathered feature rows in flight: a group of U non-zeros
//     is multiplied out of its slots while each slot is immediately refilled with the gather of the
//     non-zero U positions ahead — loads never drain at row ends;
//   * row ends are warp-uniform bookkeeping: the next 32 row-end offsets sit one per lane and are
//     broadcast by a shuffle when the stream position reaches the current row end; groups that do
//     not contain a row end (the common case) run without any per-non-zero test;
//   * (col,val) pairs arrive 32 at a time through a per-warp shared-memory slab (coalesced 256-byte
//     load, prefetched one slab ahead, read back as LDS broadcasts).
// Every output row is still produced by exactly one warp: no atomics, bit-reproducible.
constexpr int kStreamWarps = 4;
constexpr int kL2Streams = 0x100;   // stream-kernel argument bits above the GCNB_EPI_* flags
constexpr int kL2BLast = 0x200;
constexpr int kNoMedium = 0x400;   // packed kernel: row ends inside a turn take the general path only (experiments)
constexpr int kSegNnz = 768;     // target non-zeros per segment of short rows (plan time)
constexpr int kSegLong = 1024;   // a long row is cut into pieces of at most this many non-zeros

struct Seg {
  int32_t r0;    // first row
  int32_t r1;    // >= 0: one past the last row (whole rows, results go to C);
                 // <  0: a piece of row r0 whose partial sums go to partials slot (-1 - r1)
  int32_t zbeg;  // first non-zero (index into cv)
  int32_t zend;  // one past the last
};

template <typename TB, int VEC, int NV, int U, int MINB>
__global__ void __launch_bounds__(kStreamWarps * 32, MINB)
spmm_stream_kernel(const Seg* __restrict__ segs, int n_segs, const int32_t* __restrict__ rowptr,
                   const int2* __restrict__ cv, const TB* __restrict__ B, uint32_t ldb_bytes,
                   float* __restrict__ C, int64_t ldc, float* __restrict__ partials, int k,
                   const float* __restrict__ bias, int epi_l2) {
  // ring depths that divide 32 never straddle a 32-entry slab; others (U = 3: the measured per-warp cap on
  // gathers in flight) use a circular slab of two 32-entry halves
  // The slab is a ring of four 32-entry quarters filled by cp.async two quarters ahead of the one being issued
  // from: no register holds the (col,val) stream in flight (the register version stalled 5 % of the warp time
  // on its shared-memory store, profiles/r02_ae_*) and a ring turn may straddle two quarters (U = 3).
  constexpr int SL = 128;
  // bits 8.. of the last argument: L2 eviction priorities (kL2Streams: (col,val) stream and C rows are
  // evict_first; kL2BLast: gathered feature rows are evict_last), chosen by the host (spmm.cu)
  const int epi_arg = epi_l2 & 0xff;
  const uint64_t pol_s = l2_policy((epi_l2 & kL2Streams) ? 1 : 0);
  const uint64_t pol_b = l2_policy((epi_l2 & kL2BLast) ? 2 : 0);
  __shared__ alignas(16) int2 s_cv[kStreamWarps][SL];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int s = blockIdx.x * kStreamWarps + warp;
  if (s >= n_segs) return;   // warp-uniform; no block-wide barrier below
  const int4 sg = __ldg(reinterpret_cast<const int4*>(segs) + s);
  int row = sg.x;
  const bool piece = sg.y < 0;            // part of one long row: partial sums, no epilogue
  const int r1 = piece ? row + 1 : sg.y;
  const int colbase = blockIdx.y * (32 * NV * VEC);

  bool cok[NV];
  int coff[NV];
#pragma unroll
  for (int t = 0; t < NV; ++t) {
    coff[t] = colbase + (lane + t * 32) * VEC;
    cok[t] = coff[t] < k;
  }
  // NV == 1: lanes past the row's last vector point at that last vector and load without a predicate (the bytes
  // sit in the 128-byte line their quarter-warp touches anyway; their sums are never stored)
  const char* bl = reinterpret_cast<const char*>(
      B + (cok[0] ? coff[0] : (NV == 1 ? colbase + ((k - 1 - colbase) / VEC) * VEC : 0)));
  const int r1s = sg.y;

  // row ends of rows rbase .. rbase+31, one per lane
  int rbase = row;
  const int z0 = sg.z;
  const int total = sg.w - z0;
  int rp_c = piece ? sg.w : __ldg(rowptr + min(rbase + 1 + lane, r1));
  int rend = __shfl_sync(0xffffffffu, rp_c, 0) - z0;   // stream position at which `row` ends

  float acc[NV][VEC];
#pragma unroll
  for (int t = 0; t < NV; ++t)
#pragma unroll
    for (int x = 0; x < VEC; ++x) acc[t][x] = 0.f;

  // accumulate mode (C += A·B, the remote-column block of the row-partitioned SpMM): every row end is a
  // read-modify-write of a C row that nothing has touched yet, i.e. an HBM-latency load the store and
  // the next row wait for.  The rows of a segment are consecutive, so their C lines are requested into
  // L2 kPfAhead rows ahead (prefetch.global.L2: no register, no scoreboard), which turns that load into
  // an L2 hit.
  constexpr int kPfAhead = 8;
  const bool pf = (epi_arg & GCNB_EPI_ACCUM) && !piece;
  auto prefetch_c = [&](int r) {   // lanes 0..lines-1 each request one 128-byte line of C row r
    if (r < r1 && lane * 32 < k) {
      const float* p = C + (int64_t)r * ldc + lane * 32;
      asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    }
  };
  if (pf) {
#pragma unroll 1
    for (int j = 0; j < kPfAhead; ++j) prefetch_c(row + j);
  }

  auto flush = [&]() {   // write row `row`, move to the next one (warp-uniform)
    if (pf) prefetch_c(row + kPfAhead);
    // (addresses are rebuilt from the kernel parameters here, off the hot path, to keep the gather
    // loop's register count down)
    float* crow = r1s < 0 ? partials + (int64_t)(-1 - r1s) * k : C + (int64_t)row * ldc;
    const int epi = r1s < 0 ? 0 : epi_arg;
#pragma unroll
    for (int t = 0; t < NV; ++t) {
      if (cok[t]) {
        float o[VEC];
        const int nval = cols_valid<VEC>(coff[t], k);
        if (epi & GCNB_EPI_ACCUM) {
#pragma unroll
          for (int x = 0; x < VEC; ++x)
            if (x < nval) acc[t][x] += crow[coff[t] + x];
        }
#pragma unroll
        for (int x = 0; x < VEC; ++x) o[x] = epilogue(acc[t][x], bias, VEC == 8 ? min(coff[t] + x, k - 1) : coff[t] + x, epi);
        store_row_vec<VEC>(crow + coff[t], o, nval, pol_s);
      }
#pragma unroll
      for (int x = 0; x < VEC; ++x) acc[t][x] = 0.f;
    }
    ++row;
    int idx = row - rbase;
    if (idx == 32) {
      rbase = row;
      rp_c = __ldg(rowptr + min(rbase + 1 + lane, r1));
      idx = 0;
    }
    rend = __shfl_sync(0xffffffffu, rp_c, idx) - z0;
  };

  float av[U];
  RawVec<TB, VEC> bv[U][NV];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    av[u] = 0.f;
#pragma unroll
    for (int t = 0; t < NV; ++t) bv[u][t].zero();
  }
  auto consume = [&](int u) {   // acc += av[u] * (row in slot u)
#pragma unroll
    for (int t = 0; t < NV; ++t) {
      float b[VEC];
      bv[u][t].get(b);
      if constexpr (VEC % 2 == 0) {
#pragma unroll
        for (int x = 0; x < VEC; x += 2) ffma2(acc[t][x], acc[t][x + 1], av[u], b[x], b[x + 1]);
      } else {
#pragma unroll
        for (int x = 0; x < VEC; ++x) acc[t][x] = fmaf(av[u], b[x], acc[t][x]);
      }
    }
  };

  // shared-memory address of this warp's slab, made opaque so that it stays in ONE register instead
  // of being recomputed from %tid / the CTA's window base for every group
  uint32_t slab_a, slab_l;
  asm volatile("mov.u32 %0, %1;" : "=r"(slab_a) : "r"((uint32_t)__cvta_generic_to_shared(s_cv[warp])));
  asm volatile("mov.u32 %0, %1;" : "=r"(slab_l) : "r"(slab_a + 8u * (uint32_t)lane));
  const int2* src = cv + z0;
  // issue side: `nxt` = stream position of the next non-zero to gather; its slab is refilled whenever
  // nxt crosses a multiple of 32 (every earlier entry of the old slab has been issued by then)
  auto prefetch = [&](int at) {   // quarter (at / 32) % 4 <- stream positions [at, at+32), asynchronously
    if (at + lane < total)
      asm volatile("cp.async.ca.shared.global.L2::cache_hint [%0], [%1], 8, %2;"
                   ::"r"(slab_l + 8u * (uint32_t)(at & (SL - 32))), "l"(src + at + lane), "l"(pol_s) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  auto refill = [&](int at) {   // positions [at, at+32) become readable (at: multiple of 32); [at+64, at+96) starts
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncwarp();
    prefetch(at + 64);
  };
  auto gather = [&](int u, int c, int vbits) {
    av[u] = __int_as_float(vbits);
    const char* p0 = bl + (uint64_t)(uint32_t)c * ldb_bytes;
#pragma unroll
    for (int t = 0; t < NV; ++t)
      if (NV == 1 || cok[t]) bv[u][t].load(reinterpret_cast<const TB*>(p0) + t * 32 * VEC, pol_b);
  };
  auto issue = [&](int u, int at) {   // gather non-zero `at` (valid) into ring slot u
    int2 q;
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(q.x), "=r"(q.y) : "r"(slab_a + 8u * (uint32_t)(at & (SL - 1))) : "memory");
    gather(u, q.x, q.y);
  };
  auto issue2 = [&](int u, int at) {   // two consecutive non-zeros (at even) with one LDS.128 broadcast
    int4 q;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(q.x), "=r"(q.y), "=r"(q.z), "=r"(q.w)
                 : "r"(slab_a + 8u * (uint32_t)(at & (SL - 1))) : "memory");
    gather(u, q.x, q.y);
    gather(u + 1, q.z, q.w);
  };

  auto issue2_at = [&](int u, uint32_t addr) {   // the same from a ready-made slab address
    int4 q;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(q.x), "=r"(q.y), "=r"(q.z), "=r"(q.w) : "r"(addr) : "memory");
    gather(u, q.x, q.y);
    gather(u + 1, q.z, q.w);
  };

  // prologue: fill the ring
  prefetch(0);
  prefetch(32);
  refill(0);
#pragma unroll
  for (int u = 0; u < U; ++u)
    if (u < total) issue(u, u);

  for (int g = 0; g < total; g += U) {   // consume [g, g+U), issue [g+U, g+2U)
    if constexpr (32 % U == 0 && U % 2 == 0) {
      if (!(epi_l2 & kNoMedium) && (g & 31) == 0 && g + 32 <= rend && g + 32 + U <= total) {
        // a whole quarter without a row end, the turn after it complete: its 32 / U turns in one piece, slab
        // offsets constant, no per-turn tests
        const uint32_t qa = slab_a + 8u * (uint32_t)(g & (SL - 32));
        const uint32_t qb = slab_a + 8u * (uint32_t)((g + 32) & (SL - 32));
        // (rolled over the turns: fully unrolled, ptxas hoists the slab reads of all turns and spills around them)
        uint32_t sa = qa + 8u * (uint32_t)U;
#pragma unroll 1
        for (int t = 1; t < 32 / U; ++t, sa += 8u * (uint32_t)U) {
#pragma unroll
          for (int u = 0; u < U; u += 2) {
            consume(u);
            consume(u + 1);
            issue2_at(u, sa + 8u * (uint32_t)u);
          }
        }
        refill(g + 32);
#pragma unroll
        for (int u = 0; u < U; u += 2) {
          consume(u);
          consume(u + 1);
          issue2_at(u, qb + 8u * (uint32_t)u);
        }
        g += 32 - U;
        continue;
      }
    }
    const int nxt = g + U;
    if constexpr (32 % U == 0) {
      if ((nxt & 31) == 0 && nxt < total) refill(nxt);
    } else {
      // the group issued in this turn is [nxt, nxt+U): make its quarter readable when it starts one
      const int last = nxt + U - 1;
      if ((last >> 5) != ((nxt - 1) >> 5) && (last & ~31) < total) refill(last & ~31);
    }
    if (g + U <= rend && nxt + U <= total) {
      // fast path: no row end inside the group, the next group is complete
      if constexpr (U % 2 == 0) {
#pragma unroll
        for (int u = 0; u < U; u += 2) {
          consume(u);
          consume(u + 1);
          issue2(u, nxt + u);
        }
      } else {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          consume(u);
          issue(u, nxt + u);
        }
      }
    } else {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (g + u < total) {
          while (g + u == rend) flush();
          consume(u);
        }
        if (nxt + u < total) issue(u, nxt + u);
      }
    }
  }
  while (row < r1) flush();   // the last row of the segment and any trailing empty rows
}

// ---- stream class, narrow feature rows: S non-zeros per warp-wide load ---------------------------
// Measured (profiles/r01_t_*, r01_u_*): on the products-shaped graph the stream kernel above costs
// ~2.5 ms for ANY width between 24 and 64 floats and 3.6 ms for 100..128 — it is bound by the number of
// warp-wide gather requests (one per non-zero, ~6 cycles per request per SM), not by bytes.  When a
// feature row needs only L = 32/S lanes, this variant lets ONE load instruction fetch the rows of S
// consecutive non-zeros of the stream (sub-group s of L lanes takes non-zero base+s), which divides the
// request count by S.  Each sub-group keeps its own partial sums; at a row end they are folded with a
// fixed xor butterfly (deterministic) and sub-group 0 writes the row.  Everything else — segments,
// slab, ring, partial-row pieces — is as in spmm_stream_kernel.
template <typename TB, int VEC, int S, int U, int MINB>
__global__ void __launch_bounds__(kStreamWarps * 32, MINB)
spmm_stream_packed_kernel(const Seg* __restrict__ segs, int n_segs, const int32_t* __restrict__ rowptr,
                          const int2* __restrict__ cv, const TB* __restrict__ B, uint32_t ldb_bytes,
                          float* __restrict__ C, int64_t ldc, float* __restrict__ partials, int k,
                          const float* __restrict__ bias, int epi_l2) {
  // bits 8.. of the last argument: L2 eviction priorities, as in spmm_stream_kernel
  const int epi_arg = epi_l2 & 0xff;
  const uint64_t pol_s = l2_policy((epi_l2 & kL2Streams) ? 1 : 0);
  const uint64_t pol_b = l2_policy((epi_l2 & kL2BLast) ? 2 : 0);
  const bool medium = !(epi_l2 & kNoMedium);
  constexpr int L = 32 / S;     // lanes per sub-group
  constexpr int BLK = U * S;    // non-zeros per ring turn
  static_assert(S >= 2 && 32 % BLK == 0, "a ring turn must not straddle a slab");
  // (col,val) slab: four 32-entry quarters per warp, filled by cp.async two quarters ahead of the one being issued
  // from — no register holds the stream in flight and the copy has ~16 ring turns to arrive
  constexpr int SLQ = 128;
  __shared__ alignas(16) int2 s_cv[kStreamWarps][SLQ];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int s = blockIdx.x * kStreamWarps + warp;
  if (s >= n_segs) return;
  const int4 sg = __ldg(reinterpret_cast<const int4*>(segs) + s);
  const int sub = lane / L, li = lane % L;
  int row = sg.x;
  const bool piece = sg.y < 0;
  const int r1 = piece ? row + 1 : sg.y;
  const int r1s = sg.y;
  const int coff = li * VEC;
  const bool cok = coff < k;
  // lanes past the row's last vector point at that last vector: their (unpredicated) loads stay inside the
  // 128-byte line their quarter-warp touches anyway
  const char* bl = reinterpret_cast<const char*>(B + (cok ? coff : ((k - 1) / VEC) * VEC));

  int rbase = row;
  const int z0 = sg.z;
  const int total = sg.w - z0;
  int rp_c = piece ? sg.w : __ldg(rowptr + min(rbase + 1 + lane, r1));
  int rend = __shfl_sync(0xffffffffu, rp_c, 0) - z0;

  float acc[VEC];
#pragma unroll
  for (int x = 0; x < VEC; ++x) acc[x] = 0.f;

  auto flush = [&]() {
#pragma unroll
    for (int o = L; o < 32; o <<= 1)
#pragma unroll
      for (int x = 0; x < VEC; ++x) acc[x] += __shfl_xor_sync(0xffffffffu, acc[x], o);
    if (sub == 0 && cok) {
      float* crow = r1s < 0 ? partials + (int64_t)(-1 - r1s) * k : C + (int64_t)row * ldc;
      const int epi = r1s < 0 ? 0 : epi_arg;
      float o[VEC];
      const int nval = cols_valid<VEC>(coff, k);
      if (epi & GCNB_EPI_ACCUM) {
#pragma unroll
        for (int x = 0; x < VEC; ++x)
          if (x < nval) acc[x] += crow[coff + x];
      }
#pragma unroll
      for (int x = 0; x < VEC; ++x) o[x] = epilogue(acc[x], bias, VEC == 8 ? min(coff + x, k - 1) : coff + x, epi);
      store_row_vec<VEC>(crow + coff, o, nval, pol_s);
    }
#pragma unroll
    for (int x = 0; x < VEC; ++x) acc[x] = 0.f;
    ++row;
    int idx = row - rbase;
    if (idx == 32) {
      rbase = row;
      rp_c = __ldg(rowptr + min(rbase + 1 + lane, r1));
      idx = 0;
    }
    rend = __shfl_sync(0xffffffffu, rp_c, idx) - z0;
  };

  float av[U];
  RawVec<TB, VEC> bv[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    av[u] = 0.f;
    bv[u].zero();
  }
  auto consume = [&](int u) {
    float b[VEC];
    bv[u].get(b);
    if constexpr (VEC % 2 == 0) {
#pragma unroll
      for (int x = 0; x < VEC; x += 2) ffma2(acc[x], acc[x + 1], av[u], b[x], b[x + 1]);
    } else {
#pragma unroll
      for (int x = 0; x < VEC; ++x) acc[x] = fmaf(av[u], b[x], acc[x]);
    }
  };

  uint32_t slab_a, slab_l;
  asm volatile("mov.u32 %0, %1;" : "=r"(slab_a) : "r"((uint32_t)__cvta_generic_to_shared(s_cv[warp])));
  asm volatile("mov.u32 %0, %1;" : "=r"(slab_l) : "r"(slab_a + 8u * (uint32_t)lane));
  const uint32_t slab_s = slab_a + 8u * (uint32_t)sub;   // this sub-group's entry inside a slot
  const int2* src = cv + z0;
  // quarter (at / 32) % 4 <- stream positions [at, at+32), asynchronously (lanes past the end copy nothing:
  // their entries are never read as valid)
  auto prefetch = [&](int at) {
    if (at + lane < total)
      asm volatile("cp.async.ca.shared.global.L2::cache_hint [%0], [%1], 8, %2;"
                   ::"r"(slab_l + 8u * (uint32_t)(at & (SLQ - 32))), "l"(src + at + lane), "l"(pol_s) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  // make positions [at, at+32) readable (their copy was committed two refills ago) and start the copy of
  // [at+64, at+96) into the quarter that held [at-64, at-32), which nothing reads any more
  auto refill = [&](int at) {
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncwarp();
    prefetch(at + 64);
  };
  // slot u <- non-zeros at .. at+S-1 (at is a multiple of S), one per sub-group
  auto issue_full = [&](int u, int at) {
    int2 q;
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(q.x), "=r"(q.y) : "r"(slab_s + 8u * (uint32_t)(at & (SLQ - 1))) : "memory");
    av[u] = __int_as_float(q.y);
    // lanes past the row's last vector (cok false) re-read that last vector (see `bl`), so the load needs no
    // predicate; their sums are never stored
    bv[u].load(reinterpret_cast<const TB*>(bl + (uint64_t)(uint32_t)q.x * ldb_bytes), pol_b);
  };
  // the same from a slab address (of this sub-group's entry of the turn's first slot) plus a byte offset that is a
  // compile-time constant after unrolling: one address computation per turn instead of one per slot
  auto issue_slab = [&](int u, uint32_t sa, uint32_t off) {
    int2 q;
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(q.x), "=r"(q.y) : "r"(sa + off) : "memory");
    av[u] = __int_as_float(q.y);
    bv[u].load(reinterpret_cast<const TB*>(bl + (uint64_t)(uint32_t)q.x * ldb_bytes), pol_b);
  };
  auto issue_tail = [&](int u, int at) {
    int2 q;
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(q.x), "=r"(q.y) : "r"(slab_s + 8u * (uint32_t)(at & (SLQ - 1))) : "memory");
    const bool valid = at + sub < total;   // entries past the end of the stream are stale
    av[u] = valid ? __int_as_float(q.y) : 0.f;
    if (valid && cok) bv[u].load(reinterpret_cast<const TB*>(bl + (uint64_t)(uint32_t)q.x * ldb_bytes), pol_b);
    else bv[u].zero();
  };

  prefetch(0);
  prefetch(32);
  refill(0);
#pragma unroll
  for (int u = 0; u < U; ++u)
    if (u * S < total) issue_tail(u, u * S);

  for (int g = 0; g < total; g += BLK) {
    if (medium && (g & 31) == 0 && g + 32 <= rend && g + 32 + BLK <= total) {
      // a whole slab without a row end, and the turn after it complete: its 32 / BLK turns in one piece — slab
      // offsets are constants, no per-turn tests (about half of all non-zeros at 50 per row)
      const uint32_t qa = slab_s + 8u * (uint32_t)(g & (SLQ - 32));              // quarter of [g, g+32)
      const uint32_t qb = slab_s + 8u * (uint32_t)((g + 32) & (SLQ - 32));       // the one after it
      // (a rolled loop over the turns of the quarter: fully unrolled, ptxas hoists the slab reads of all turns
      // and spills ring state around them)
      uint32_t sa = qa + 8u * (uint32_t)BLK;
#pragma unroll 1
      for (int t = 1; t < 32 / BLK; ++t, sa += 8u * (uint32_t)BLK) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          consume(u);
          issue_slab(u, sa, 8u * (uint32_t)(u * S));
        }
      }
      refill(g + 32);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        consume(u);
        issue_slab(u, qb, 8u * (uint32_t)(u * S));
      }
      g += 32 - BLK;
      continue;
    }
    const int nxt = g + BLK;
    if ((nxt & 31) == 0 && nxt < total) refill(nxt);
    if (nxt <= rend && nxt + BLK <= total) {
      // fast path: no row end inside this turn, the next turn is complete
      const uint32_t sa = slab_s + 8u * (uint32_t)(nxt & (SLQ - 1));
#pragma unroll
      for (int u = 0; u < U; ++u) {
        consume(u);
        issue_slab(u, sa, 8u * (uint32_t)(u * S));
      }
    } else if (medium && nxt + BLK <= total) {
      // a row ends inside this turn, the next turn is complete (every turn but a segment's last two): slots the
      // row end does not touch are consumed whole, the next turn is issued without validity tests
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int base = g + u * S;
        if (rend >= base + S) {
          consume(u);
        } else {
          const int pos = base + sub;
          int done = base;
          while (true) {
            while (done == rend) flush();
            const int lim = min(rend, base + S);
            if (pos >= done && pos < lim) consume(u);
            done = lim;
            if (done >= base + S) break;
          }
        }
        issue_full(u, nxt + u * S);
      }
    } else {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int base = g + u * S;          // slot u holds stream positions [base, base+S)
        if (base < total) {
          const int pos = base + sub;
          const int stop = min(base + S, total);
          int done = base;
          while (true) {
            while (done == rend) flush();      // the current row ends here (also skips empty rows)
            const int lim = min(rend, stop);
            if (pos >= done && pos < lim) consume(u);
            done = lim;
            if (done >= stop) break;
          }
        }
        if (nxt + u * S < total) issue_tail(u, nxt + u * S);
      }
    }
  }
  while (row < r1) flush();
}

// ---- panel class: 32 consecutive rows walk their columns in step, so that a gathered feature row is
// ---- re-used by the other rows of the panel out of L1 ---------------------------------------------
// The one idea the reference's tile format exists for (8-row column-ordered panels, tile.cu:11-103,
// consumed per segment in flexspmm.cu:59-92): rows that are processed TOGETHER and reference the same
// column should fetch that feature row from L2 once.  The stream kernel above cannot get that re-use: its
// warps run through unrelated row segments at unrelated times, every non-zero is a 512-byte L2->SM
// transfer, and on graphs with mean degree in the hundreds (Reddit-shaped) the 64 B/clk/SM L2 port is
// the bound.  Here a CTA of 32 warps owns a PANEL of 32 consecutive rows, one row per warp, accumulators
// in registers, no atomics.  At plan time (spmm.cu build_panels) the panel's column range is cut into
// blocks of ~kPanelTileNnz non-zeros of the panel (a few hundred distinct 512-byte feature rows: an
// L1-sized working set) and every row gets the offsets of its non-zeros at those cuts.  All 32 warps
// process block t before any of them goes beyond block t+1 (two mbarriers, one block of drift allowed),
// so a feature row gathered by one warp is an L1 hit for the other rows of the panel that reference it
// inside the same block: nnz / distinct-columns-per-block L2 transfers are saved (1.8x on the
// Reddit-shaped graph in degree order, 20x inside its hub rows).  Everything else is the stream kernel:
// (col,val) slab, register ring of U gathers in flight, fma.rn.f32x2.
// Heavy panels are cut along their block list into ITEMS of <= ~32 k non-zeros (one CTA each) whose
// partial rows meet in spmm_reduce_kernel in fixed order: still no atomics, bit-reproducible.
constexpr int kPanelRows = 32;

struct PanelItem {
  int32_t row0;      // first row of the panel
  int32_t a, b;      // blocks [a, b) of the panel
  int32_t slot;      // < 0: the item is the whole panel, rows go to C; else first partial row of the item
  int32_t n_items;   // items of this panel (partial row of panel row r = slot + r * n_items)
  int32_t sub_off;   // offset of the panel's boundary table in subptr
  int32_t nb;        // blocks in the panel (the table has kPanelRows rows of nb + 1 offsets)
  int32_t pad;
};

template <typename TB, int VEC, int U>
__global__ void __launch_bounds__(kPanelRows * 32, 1)
spmm_panel_kernel(const PanelItem* __restrict__ items, const int32_t* __restrict__ subptr,
                  const int2* __restrict__ cv, const TB* __restrict__ B, uint32_t ldb_bytes,
                  float* __restrict__ C, int64_t ldc, float* __restrict__ partials, int m, int k,
                  const float* __restrict__ bias, int epi_l2) {
  static_assert(32 % U == 0, "ring size must divide the slab");
  __shared__ alignas(16) int2 s_cv[kPanelRows][32];
  __shared__ alignas(8) uint64_t s_bar[2];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int epi_arg = epi_l2 & 0xff;
  const uint64_t pol_s = l2_policy((epi_l2 & kL2Streams) ? 1 : 0);
  const uint64_t pol_b = l2_policy((epi_l2 & kL2BLast) ? 2 : 0);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_bar[0])), "r"(kPanelRows));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_bar[1])), "r"(kPanelRows));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(&s_bar[0]);

  const PanelItem it = items[blockIdx.x];
  const int row = it.row0 + warp;
  const int nblk = it.b - it.a;
  const int32_t* bp = subptr + it.sub_off + (int64_t)warp * (it.nb + 1) + it.a;   // bp[t] = start of block a+t
  const int z0 = __ldg(bp);
  const int total = __ldg(bp + nblk) - z0;
  const int coff = (blockIdx.y * 32 + lane) * VEC;
  const bool cok = coff < k;
  const char* bl = reinterpret_cast<const char*>(B + (cok ? coff : 0));

  // ends of blocks tb .. tb+31 (stream positions), one per lane
  int tb = 0, t = 0;
  int bnd_c = __ldg(bp + min(1 + lane, nblk)) - z0;
  int rend = __shfl_sync(0xffffffffu, bnd_c, 0);

  float acc[VEC];
#pragma unroll
  for (int x = 0; x < VEC; ++x) acc[x] = 0.f;

  auto boundary = [&]() {   // the stream position reached the end of block t (warp-uniform)
    if (t == nblk - 1) {
      if (row < m && cok) {
        float* crow = it.slot < 0 ? C + (int64_t)row * ldc
                                  : partials + ((int64_t)it.slot + (int64_t)warp * it.n_items) * k;
        const int epi = it.slot < 0 ? epi_arg : 0;
        float o[VEC];
#pragma unroll
        for (int x = 0; x < VEC; ++x) o[x] = epilogue(acc[x], bias, coff + x, epi);
        store_vec<VEC>(crow + coff, o, pol_s);
      }
    } else {
      // arrive on the barrier of block t; before starting block t+1 the whole panel must have finished
      // block t-1 (one block of drift)
      __syncwarp();
      if (lane == 0)
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar0 + 8u * (uint32_t)(t & 1)) : "memory");
      if (t >= 1) {
        const uint32_t bar = bar0 + 8u * (uint32_t)((t - 1) & 1), parity = (uint32_t)(((t - 1) >> 1) & 1);
        uint32_t done = 0;
        while (!done)
          asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                       : "=r"(done) : "r"(bar), "r"(parity) : "memory");
      }
    }
    ++t;
    int idx = t - tb;
    if (idx == 32) {
      tb = t;
      bnd_c = __ldg(bp + min(1 + tb + lane, nblk)) - z0;
      idx = 0;
    }
    rend = __shfl_sync(0xffffffffu, bnd_c, idx);
  };

  float av[U];
  RawVec<TB, VEC> bv[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    av[u] = 0.f;
    bv[u].zero();
  }
  auto consume = [&](int u) {
    float b[VEC];
    bv[u].get(b);
    if constexpr (VEC % 2 == 0) {
#pragma unroll
      for (int x = 0; x < VEC; x += 2) ffma2(acc[x], acc[x + 1], av[u], b[x], b[x + 1]);
    } else {
#pragma unroll
      for (int x = 0; x < VEC; ++x) acc[x] = fmaf(av[u], b[x], acc[x]);
    }
  };

  uint32_t slab_a, slab_l;
  asm volatile("mov.u32 %0, %1;" : "=r"(slab_a) : "r"((uint32_t)__cvta_generic_to_shared(s_cv[warp])));
  asm volatile("mov.u32 %0, %1;" : "=r"(slab_l) : "r"(slab_a + 8u * (uint32_t)lane));
  const int2* src = cv + z0;
  int2 e = make_int2(0, 0);
  if (lane < total) e = ld_stream_int2(src + lane, pol_s);
  auto refill = [&](int at) {
    __syncwarp();
    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(slab_l), "r"(e.x), "r"(e.y) : "memory");
    __syncwarp();
    if (at + 32 + lane < total) e = ld_stream_int2(src + at + 32 + lane, pol_s);
  };
  auto gather = [&](int u, int c, int vbits) {
    av[u] = __int_as_float(vbits);
    if (cok) bv[u].load(reinterpret_cast<const TB*>(bl + (uint64_t)(uint32_t)c * ldb_bytes), pol_b);
  };
  auto issue = [&](int u, int at) {
    int2 q;
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(q.x), "=r"(q.y) : "r"(slab_a + 8u * (uint32_t)(at & 31)) : "memory");
    gather(u, q.x, q.y);
  };
  auto issue2 = [&](int u, int at) {
    int4 q;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(q.x), "=r"(q.y), "=r"(q.z), "=r"(q.w)
                 : "r"(slab_a + 8u * (uint32_t)(at & 31)) : "memory");
    gather(u, q.x, q.y);
    gather(u + 1, q.z, q.w);
  };

  if (total > 0) {
    refill(0);
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (u < total) issue(u, u);
  }
  for (int g = 0; g < total; g += U) {
    const int nxt = g + U;
    if ((nxt & 31) == 0 && nxt < total) refill(nxt);
    if (g + U <= rend && nxt + U <= total) {
      if constexpr (U % 2 == 0) {
#pragma unroll
        for (int u = 0; u < U; u += 2) {
          consume(u);
          consume(u + 1);
          issue2(u, nxt + u);
        }
      } else {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          consume(u);
          issue(u, nxt + u);
        }
      }
    } else {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (g + u < total) {
          while (g + u == rend) boundary();
          consume(u);
        }
        if (nxt + u < total) issue(u, nxt + u);
      }
    }
  }
  while (t < nblk) boundary();   // the last block of the item and any trailing empty ones
}

// ---- block class -------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

template <typename TB, int VEC, int G, int NV>
__global__ void __launch_bounds__(kBlockWarps * 32)
spmm_block_kernel(const WorkItem* __restrict__ items, const int2* __restrict__ cv,
                  const TB* __restrict__ B, int64_t ldb, float* __restrict__ C, int64_t ldc,
                  float* __restrict__ partials, int k, const float* __restrict__ bias, int epi) {
  constexpr int RPW = 32 / G;  // non-zeros handled per warp per step
  constexpr int W = G * NV * VEC;
  constexpr int U = unroll_for<NV, VEC>();
  __shared__ alignas(16) int2 s_cv[kChunk + 2];
  __shared__ alignas(16) float s_red[kBlockWarps][W];
  __shared__ alignas(8) uint64_t s_bar;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int lig = lane & (G - 1), grp = lane / G;
  const int colbase = blockIdx.y * W;
  const WorkItem it = items[blockIdx.x];

  // One TMA bulk copy brings the item's (col,val) tile into shared memory.  Source and size must
  // be 16-byte aligned: start one pair early when beg is odd (cv is padded by 2 pairs).
  const int abeg = it.beg & ~1;
  const uint32_t bytes = (uint32_t)(((it.end - abeg) * 8 + 15) & ~15);
  const uint32_t bar = smem_u32(&s_bar);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (tid == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
                 : "memory");
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(smem_u32(s_cv)), "l"(cv + abeg), "r"(bytes), "r"(bar)
        : "memory");
  }
  {
    uint32_t done = 0;
    while (!done) {
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\t"
          "selp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(done)
          : "r"(bar)
          : "memory");
    }
  }

  int coff[NV];
  bool cok[NV];
#pragma unroll
  for (int t = 0; t < NV; ++t) {
    coff[t] = colbase + (lig + t * G) * VEC;
    cok[t] = coff[t] < k;
  }
  float acc[NV][VEC];
#pragma unroll
  for (int t = 0; t < NV; ++t)
#pragma unroll
    for (int x = 0; x < VEC; ++x) acc[t][x] = 0.f;

  const int off = it.beg - abeg;
  const int n = it.end - it.beg;
  constexpr int stride = kBlockWarps * RPW;
  const int j_first = warp * RPW + grp;
  for (int j0 = 0; j0 < n; j0 += stride * U) {
    float bv[U][NV][VEC];
    float av[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int j = j0 + u * stride + j_first;
      const bool p = j < n;
      const int2 e = p ? s_cv[off + j] : make_int2(0, 0);
      av[u] = p ? __int_as_float(e.y) : 0.f;
      const TB* bp = B + (int64_t)e.x * ldb;
#pragma unroll
      for (int t = 0; t < NV; ++t) {
        if (p && cok[t]) {
          VecLoad<TB, VEC>::load(bp + coff[t], bv[u][t]);
        } else {
#pragma unroll
          for (int x = 0; x < VEC; ++x) bv[u][t][x] = 0.f;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int t = 0; t < NV; ++t)
#pragma unroll
        for (int x = 0; x < VEC; ++x) acc[t][x] = fmaf(av[u], bv[u][t][x], acc[t][x]);
  }

  // lane groups of one warp hold different non-zeros of the same row: fold them
  if constexpr (G < 32) {
#pragma unroll
    for (int o = G; o < 32; o <<= 1)
#pragma unroll
      for (int t = 0; t < NV; ++t)
#pragma unroll
        for (int x = 0; x < VEC; ++x) acc[t][x] += __shfl_xor_sync(0xffffffffu, acc[t][x], o);
  }
  if (grp == 0) {
#pragma unroll
    for (int t = 0; t < NV; ++t)
#pragma unroll
      for (int x = 0; x < VEC; ++x) s_red[warp][(lig + t * G) * VEC + x] = acc[t][x];
  }
  __syncthreads();
  for (int i = tid; i < W; i += kBlockWarps * 32) {
    const int col = colbase + i;
    if (col < k) {
      float s = s_red[0][i];
#pragma unroll
      for (int w = 1; w < kBlockWarps; ++w) s += s_red[w][i];
      if (it.slot < 0) {
        if (epi & GCNB_EPI_ACCUM) s += C[(int64_t)it.row * ldc + col];
        __stcs(C + (int64_t)it.row * ldc + col, epilogue(s, bias, col, epi));
      }
      else
        partials[(int64_t)it.slot * k + col] = s;
    }
  }
}

// rows that were cut into several items: sum their partial rows in item order
__global__ void __launch_bounds__(128)
spmm_reduce_kernel(const int32_t* __restrict__ split_rows, const int32_t* __restrict__ split_ptr,
                   const float* __restrict__ partials, float* __restrict__ C, int64_t ldc, int k,
                   const float* __restrict__ bias, int epi) {
  const int s = blockIdx.x;
  const int64_t row = split_rows[s];
  const int p0 = split_ptr[s], p1 = split_ptr[s + 1];
  for (int col = threadIdx.x; col < k; col += blockDim.x) {
    float v = (epi & GCNB_EPI_ACCUM) ? C[row * ldc + col] : 0.f;
    for (int p = p0; p < p1; ++p) v += partials[(int64_t)p * k + col];
    __stcs(C + row * ldc + col, epilogue(v, bias, col, epi));
  }
}

// (col[], val[]) -> interleaved pairs, once at plan creation
__global__ void interleave_kernel(const int32_t* __restrict__ col, const float* __restrict__ val,
                                  int2* __restrict__ cv, int64_t nnz) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nnz) cv[i] = make_int2(col[i], __float_as_int(val[i]));
}

// out[r,:] = X[vomp[r],:]   (replaces flexspmm_v9_permuteX + put_back, permutate.cu:3-39)
__global__ void __launch_bounds__(256)
permute_rows_kernel(const float* __restrict__ X, const int32_t* __restrict__ vomp,
                    float* __restrict__ out, int64_t n, int k, int vec4) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= n) return;
  const int64_t src = __ldg(vomp + row);
  if (vec4) {
    const float4* s = reinterpret_cast<const float4*>(X + src * k);
    float4* d = reinterpret_cast<float4*>(out + row * k);
    for (int i = lane; i < (k >> 2); i += 32) __stcs(d + i, __ldg(s + i));
  } else {
    for (int i = lane; i < k; i += 32) out[row * k + i] = __ldg(X + src * k + i);
  }
}

}  // namespace gcnb
