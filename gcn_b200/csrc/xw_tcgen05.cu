// Dense feature transform on the 5th-generation tensor cores:  D = op(X) · op(W) (+bias)(ReLU).
//
// Replaces torch.mm(input, self.weight) of the reference layer (pygcn/gcn6.py:131, :191;
// gcn5.py:88, :136) and the two products autograd derives from it.  This is the only GEMM-shaped
// work on the GCN path; everything else is the SpMM.
//
// Shape of the problem: M (vertices) is huge, N and K (feature widths) are small, so the kernel is
// bound by streaming X in and D out, not by the tensor pipe.  Design:
//
//   * CTA tile 128 x BN (BN = 16..256), K consumed in blocks of 16.  One elected thread issues
//     tcgen05.mma (cta_group::1, M = 128, N = BN) with both operands in shared memory in the
//     canonical K-major / no-swizzle core-matrix layout; the fp32 accumulator lives in TMEM
//     (BN columns x 128 lanes) and is read back once with tcgen05.ld for the fused bias / ReLU
//     epilogue.
//   * fp32 parity mode (precision 0): every fp32 operand value x is split on the fly into
//     hi = tf32(x) and lo = tf32(x - hi); three kind::tf32 passes hi·hi + hi·lo + lo·hi recover
//     fp32-level accuracy (~1e-6 relative).  precision 1 issues the hi·hi pass only; precision 2
//     converts to bf16 and issues kind::f16.
//   * A operand (the tall matrix): raw fp32 rows arrive through a 4-deep cp.async ring (zero-filled
//     at the edges), each thread then converts ITS OWN chunks and stores them into the canonical
//     layout — no barrier is needed between the async copy and the conversion.
//   * B operand (the small matrix, shared by every CTA): split and laid out ONCE by a prep kernel
//     into a global "image" that is byte-for-byte the shared-memory layout of each (n-tile,
//     k-block); the main kernel brings a block in with one TMA bulk copy (cp.async.bulk +
//     mbarrier complete_tx).  The prep kernel also absorbs trans_w and any leading dimension.
//   * Two canonical stages: the MMAs of block i run asynchronously (tcgen05.commit -> mbarrier)
//     while the CTA converts block i+1.  Split-K (grid.z) with a deterministic two-pass reduction
//     covers the tall-K product dW = X^T·dS.
//   * Tall-K variant (BRAW): in dW = X^T·dS BOTH operands are vertex-sized, so an image of the second
//     one costs a write and a re-read of 8 bytes per element of it (2.5 GB for the products-shaped
//     layer 1).  There the B operand takes the same road as the A operand: raw rows through a second
//     cp.async ring, split and transposed into the canonical layout by the thread that fetched them.
#include <cuda_bf16.h>

#include <stdlib.h>

#include <algorithm>

#include <atomic>

#include "common.h"

namespace gcnb {
namespace xw {

constexpr int BM = 128;       // rows per CTA tile == UMMA M
constexpr int BK = 16;        // k per stage
constexpr int THREADS = 128;  // 4 warps: warp w owns TMEM lanes 32w..32w+31 in the epilogue
constexpr int RAW_STAGES = 4; // cp.async ring depth for the raw fp32 A tile
constexpr int A_LBO = BM * 16 + 16;        // bytes between 16-byte k-chunks of the A tile (+16: bank spread)
constexpr int A_PART = (BK / 4) * A_LBO;   // bytes of one A part (hi or lo), sized for the tf32 case
constexpr int RAW_BYTES = BM * BK * 4;

__host__ __device__ constexpr int b_part_bytes(int bn) { return bn * BK * 4; }   // one B part (tf32 sizing)
// B operand converted in the kernel (BRAW): k-chunk pitch padded like the A tile's, raw ring of its own
__host__ __device__ constexpr int b_lbo(int bn, bool braw) { return braw ? bn * 16 + 16 : bn * 16; }
__host__ __device__ constexpr int b_part(int bn, bool braw) { return braw ? (BK / 4) * b_lbo(bn, true) : b_part_bytes(bn); }
__host__ __device__ constexpr int b_raw_bytes(int bn) { return bn * BK * 4; }
__host__ __device__ constexpr int stage_bytes(int bn, bool braw = false) { return 2 * A_PART + 2 * b_part(bn, braw); }
__host__ __device__ constexpr int smem_bytes(int bn, bool braw = false) {
  return 2 * stage_bytes(bn, braw) + RAW_STAGES * RAW_BYTES + (braw ? RAW_STAGES * b_raw_bytes(bn) : 0) + 64;
}

// byte offset of element (row, k) inside one operand part; K-major core matrices of 8 rows x 16 B
template <int EPC, int EB>   // elements per 16-byte chunk, bytes per element
__host__ __device__ __forceinline__ int canon_off(int row, int k, int lbo) {
  return (k / EPC) * lbo + (row >> 3) * 128 + (row & 7) * 16 + (k % EPC) * EB;
}

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint32_t to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}

// shared-memory matrix descriptor: K-major, SWIZZLE_NONE, version 1 (Blackwell)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3fff);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}

// instruction descriptor: D fp32, A/B format fmt (2 = TF32, 1 = BF16), both K-major, M = 128
__host__ __device__ constexpr uint32_t make_idesc(int fmt, int n) {
  return (1u << 4) | ((uint32_t)fmt << 7) | ((uint32_t)fmt << 10) | ((uint32_t)(n >> 3) << 17) |
         ((uint32_t)(BM >> 4) << 24);
}

template <bool TF32>
__device__ __forceinline__ void mma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  if constexpr (TF32) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
  }
}

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  }
}

// 16-byte async copy global -> shared, `bytes` (0..16) taken from the source, the rest zero-filled
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src, int bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
}

struct Params {
  const float* X; int64_t ldx; int trans_x;   // op(X) is M x K
  const uint8_t* w_img;                       // prepared B operand image (null in the BRAW variant)
  const float* W; int64_t ldw; int w_vec;     // BRAW: raw K x N row-major second operand
  float* D; int64_t ldd;                      // output (or partials when splits > 1)
  int64_t M; int N; int K;
  int nkb;                                    // k-blocks in total
  int ntiles;                                 // n-tiles
  int kb_per_split;
  const float* bias; int epi;
  int x_vec;                                  // 16-byte cp.async allowed (alignment of X and ldx)
  int d_vec;                                  // 16-byte stores allowed
};


// ---- B operand image ---------------------------------------------------------------------------
// image[(nt * nkb + kb)] = { part hi | part lo }, each part the canonical layout of a BN x BK tile
// (rows = n, K-major).  MODE 0/1: tf32 bit patterns; MODE 2: bf16.  One thread per (n, k).
template <int MODE>
__global__ void __launch_bounds__(256)
prep_w_kernel(const float* __restrict__ W, int64_t ldw, int trans_w, int N, int K, int bn, int nkb,
              int ntiles, uint8_t* __restrict__ img) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t kpad = (int64_t)nkb * BK;
  const int64_t total = (int64_t)ntiles * bn * kpad;
  if (idx >= total) return;
  const int64_t npad = (int64_t)ntiles * bn;
  int k, n;   // fastest index follows the contiguous direction of the stored matrix
  if (trans_w) { k = (int)(idx % kpad); n = (int)(idx / kpad); }
  else         { n = (int)(idx % npad); k = (int)(idx / npad); }
  float v = 0.f;
  if (n < N && k < K) v = trans_w ? W[(int64_t)n * ldw + k] : W[(int64_t)k * ldw + n];
  const int nt = n / bn, row = n % bn, kb = k / BK, kk = k % BK;
  uint8_t* tile = img + ((int64_t)nt * nkb + kb) * (2 * b_part_bytes(bn));
  if constexpr (MODE == 2) {
    const int off = canon_off<8, 2>(row, kk, bn * 16);
    *reinterpret_cast<__nv_bfloat16*>(tile + off) = __float2bfloat16_rn(v);
  } else {
    const int off = canon_off<4, 4>(row, kk, bn * 16);
    const uint32_t hi = to_tf32(v);
    *reinterpret_cast<uint32_t*>(tile + off) = hi;
    if constexpr (MODE == 0)
      *reinterpret_cast<uint32_t*>(tile + b_part_bytes(bn) + off) = to_tf32(v - __uint_as_float(hi));
  }
}

// ---- main kernel ---------------------------------------------------------------------------------
template <int BN, int MODE, bool BRAW>
__global__ void __launch_bounds__(THREADS)
xw_kernel(const Params p) {
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr bool TF32 = MODE != 2;
  constexpr int B_LBO = b_lbo(BN, BRAW);
  constexpr int B_PART = b_part(BN, BRAW);
  constexpr int STAGE = stage_bytes(BN, BRAW);
  constexpr int B_RAW = b_raw_bytes(BN);
  constexpr int NBC = BN * 4;                              // 16-byte chunks of one raw B tile (BK x BN)
  constexpr int B_BYTES = (MODE == 0 ? 2 : 1) * (TF32 ? B_PART : B_PART / 2);  // bytes the MMAs read
  constexpr uint32_t IDESC = make_idesc(TF32 ? 2 : 1, BN);
  constexpr int TMEM_COLS = BN <= 32 ? 32 : (BN <= 64 ? 64 : (BN <= 128 ? 128 : 256));

  uint8_t* raw = smem + 2 * STAGE;                         // cp.async ring, thread-private slots
  uint8_t* braw = raw + RAW_STAGES * RAW_BYTES;            // BRAW: ring of raw B tiles
  uint64_t* bars = reinterpret_cast<uint64_t*>(braw + (BRAW ? RAW_STAGES * B_RAW : 0));  // [0,1] mma done, [2,3] B landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // n-tile fastest: the CTAs that share an X tile run together, so X is fetched from HBM once
  const int nt = (int)(blockIdx.x % p.ntiles);
  const int64_t m0 = (int64_t)(blockIdx.x / p.ntiles) * BM;
  const int kb0 = blockIdx.y * p.kb_per_split;
  const int kb1 = min(p.nkb, kb0 + p.kb_per_split);
  const int nk = kb1 - kb0;                                 // >= 1 by construction of the grid

  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(bars + i)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_addr(tmem_slot)), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_d = *tmem_slot;

  // ---- raw A tile: each thread owns 4 chunks (16 bytes = 4 consecutive elements of the contiguous
  // direction) per k-block.  K-contiguous source: chunk c -> row c/4, k-chunk c%4.
  // M-contiguous source (trans_x): chunk c -> k = c%16, m-chunk c/16 (conflict-free transposing stores).
  auto issue_raw = [&](int kb, int slot) {
    const int k0 = kb * BK;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = tid + j * THREADS;
      const uint32_t dst = smem_addr(raw + slot * RAW_BYTES + c * 16);
      int64_t r, q;       // element (row r of op(X), k index q), 4 elements along the contiguous direction
      const float* src;
      int valid;          // how many of the 4 elements are inside the matrix
      if (!p.trans_x) {
        r = m0 + (c >> 2); q = k0 + (c & 3) * 4;
        valid = (r < p.M) ? max(0, min(4, p.K - (int)q)) : 0;
        src = p.X + r * p.ldx + q;
      } else {
        q = k0 + (c & 15); r = m0 + (c >> 4) * 4;
        valid = (q < p.K) ? (int)max((int64_t)0, min((int64_t)4, p.M - r)) : 0;
        src = p.X + q * p.ldx + r;
      }
      if (valid == 0) {
        cp_async16(dst, p.X, 0);                       // pure zero fill
      } else if (p.x_vec) {
        cp_async16(dst, src, valid * 4);
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) cp_async4(dst + 4 * e, e < valid ? src + e : p.X, e < valid ? 4 : 0);
      }
    }
    if constexpr (BRAW) {
      // raw B tile, N-contiguous source: chunk c -> k = c%16, n-chunk c/16 (as the trans_x A tile)
#pragma unroll
      for (int j = 0; j < (NBC + THREADS - 1) / THREADS; ++j) {
        const int c = tid + j * THREADS;
        if (NBC % THREADS != 0 && c >= NBC) break;
        const uint32_t dst = smem_addr(braw + slot * B_RAW + c * 16);
        const int q = k0 + (c & 15);
        const int n0 = nt * BN + (c >> 4) * 4;
        const int valid = (q < p.K) ? max(0, min(4, p.N - n0)) : 0;
        const float* src = p.W + (int64_t)q * p.ldw + n0;
        if (valid == 0) {
          cp_async16(dst, p.W, 0);
        } else if (p.w_vec) {
          cp_async16(dst, src, valid * 4);
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e) cp_async4(dst + 4 * e, e < valid ? src + e : p.W, e < valid ? 4 : 0);
        }
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

#pragma unroll
  for (int s = 0; s < RAW_STAGES - 1; ++s) {
    if (s < nk) issue_raw(kb0 + s, s);
    else asm volatile("cp.async.commit_group;" ::: "memory");
  }

  for (int i = 0; i < nk; ++i) {
    const int s = i & 1;
    uint8_t* stage = smem + s * STAGE;
    // the MMAs that read this canonical stage two iterations ago must have finished
    if (i >= 2) {
      mbar_wait(smem_addr(bars + s), ((i >> 1) - 1) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
    if (!BRAW && tid == 0) {   // B block: one TMA bulk copy of the prepared image
      const uint8_t* src = p.w_img + ((int64_t)nt * p.nkb + kb0 + i) * (2 * B_PART);
      const uint32_t bar = smem_addr(bars + 2 + s);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(B_BYTES) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(smem_addr(stage + 2 * A_PART)), "l"(src), "r"(B_BYTES), "r"(bar) : "memory");
    }
    asm volatile("cp.async.wait_group %0;" ::"n"(RAW_STAGES - 2) : "memory");
    // convert this thread's own chunks into the canonical A tile(s)
    const uint8_t* rslot = raw + (i % RAW_STAGES) * RAW_BYTES;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = tid + j * THREADS;
      const float4 v = *reinterpret_cast<const float4*>(rslot + c * 16);
      const float x[4] = {v.x, v.y, v.z, v.w};
      if (!p.trans_x) {
        const int row = c >> 2, k = (c & 3) * 4;
        if constexpr (TF32) {
          uint32_t hi[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) hi[e] = to_tf32(x[e]);
          const int off = canon_off<4, 4>(row, k, A_LBO);
          *reinterpret_cast<uint4*>(stage + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
          if constexpr (MODE == 0) {
            uint32_t lo[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) lo[e] = to_tf32(x[e] - __uint_as_float(hi[e]));
            *reinterpret_cast<uint4*>(stage + A_PART + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
          }
        } else {
          const __nv_bfloat162 a = __floats2bfloat162_rn(x[0], x[1]), b = __floats2bfloat162_rn(x[2], x[3]);
          const int off = canon_off<8, 2>(row, k, A_LBO);
          *reinterpret_cast<uint2*>(stage + off) =
              make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
        }
      } else {
        const int k = c & 15, row0 = (c >> 4) * 4;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          if constexpr (TF32) {
            const int off = canon_off<4, 4>(row0 + e, k, A_LBO);
            const uint32_t hi = to_tf32(x[e]);
            *reinterpret_cast<uint32_t*>(stage + off) = hi;
            if constexpr (MODE == 0)
              *reinterpret_cast<uint32_t*>(stage + A_PART + off) = to_tf32(x[e] - __uint_as_float(hi));
          } else {
            *reinterpret_cast<__nv_bfloat16*>(stage + canon_off<8, 2>(row0 + e, k, A_LBO)) =
                __float2bfloat16_rn(x[e]);
          }
        }
      }
    }
    if constexpr (BRAW) {   // the same for this thread's chunks of the raw B tile (rows of the operand = n)
      const uint8_t* bslot = braw + (i % RAW_STAGES) * B_RAW;
      uint8_t* bst = stage + 2 * A_PART;
#pragma unroll
      for (int j = 0; j < (NBC + THREADS - 1) / THREADS; ++j) {
        const int c = tid + j * THREADS;
        if (NBC % THREADS != 0 && c >= NBC) break;
        const float4 v = *reinterpret_cast<const float4*>(bslot + c * 16);
        const float x[4] = {v.x, v.y, v.z, v.w};
        const int k = c & 15, row0 = (c >> 4) * 4;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          if constexpr (TF32) {
            const int off = canon_off<4, 4>(row0 + e, k, B_LBO);
            const uint32_t hi = to_tf32(x[e]);
            *reinterpret_cast<uint32_t*>(bst + off) = hi;
            if constexpr (MODE == 0)
              *reinterpret_cast<uint32_t*>(bst + B_PART + off) = to_tf32(x[e] - __uint_as_float(hi));
          } else {
            *reinterpret_cast<__nv_bfloat16*>(bst + canon_off<8, 2>(row0 + e, k, B_LBO)) =
                __float2bfloat16_rn(x[e]);
          }
        }
      }
    }
    // refill the ring slot consumed one iteration ago (this thread's own slot: no hazard)
    if (i + RAW_STAGES - 1 < nk) issue_raw(kb0 + i + RAW_STAGES - 1, (i + RAW_STAGES - 1) % RAW_STAGES);
    else asm volatile("cp.async.commit_group;" ::: "memory");
    // generic-proxy stores -> visible to the tensor core's async proxy
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (tid == 0) {
      if constexpr (!BRAW) mbar_wait(smem_addr(bars + 2 + s), (i >> 1) & 1);    // B block landed
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t a_hi = smem_addr(stage), a_lo = a_hi + A_PART;
      const uint32_t b_hi = smem_addr(stage + 2 * A_PART), b_lo = b_hi + (TF32 ? B_PART : 0);
      constexpr int NMMA = TF32 ? BK / 8 : BK / 16;        // MMAs per part per stage (K = 32 bytes each)
#pragma unroll
      for (int kk = 0; kk < NMMA; ++kk) {
        const uint32_t ao = kk * 2 * A_LBO, bo = kk * 2 * B_LBO;   // two 16-byte k-chunks per MMA
        const uint32_t first = (i > 0 || kk > 0) ? 1u : 0u;
        mma<TF32>(tmem_d, make_desc(a_hi + ao, A_LBO, 128), make_desc(b_hi + bo, B_LBO, 128), IDESC, first);
        if constexpr (MODE == 0) {
          mma<TF32>(tmem_d, make_desc(a_hi + ao, A_LBO, 128), make_desc(b_lo + bo, B_LBO, 128), IDESC, 1u);
          mma<TF32>(tmem_d, make_desc(a_lo + ao, A_LBO, 128), make_desc(b_hi + bo, B_LBO, 128), IDESC, 1u);
        }
      }
      // arrives on the mbarrier once every MMA issued so far has completed
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                   ::"r"(smem_addr(bars + s)) : "memory");
    }
  }

  // ---- epilogue: TMEM -> registers -> bias / ReLU -> global ------------------------------------
  {
    const int last = nk - 1;
    mbar_wait(smem_addr(bars + (last & 1)), (last >> 1) & 1);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }
  const int64_t row = m0 + warp * 32 + lane;
  const int ncol0 = nt * BN;
  float* drow = p.D + ((int64_t)blockIdx.y * p.M + row) * p.ldd;   // split z writes its own partial matrix
#pragma unroll 1
  for (int c0 = 0; c0 < BN; c0 += 16) {
    if (ncol0 + c0 >= p.N) break;   // warp-uniform
    uint32_t r[16];
    const uint32_t taddr = tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0;
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    if (row < p.M) {
      float o[16];
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        const int col = ncol0 + c0 + e;
        float x = __uint_as_float(r[e]);
        if (col < p.N) {
          if (p.epi & GCNB_EPI_BIAS) x += __ldg(p.bias + col);
          if (p.epi & GCNB_EPI_RELU) x = fmaxf(x, 0.f);
        }
        o[e] = x;
      }
      if (p.d_vec && ncol0 + c0 + 16 <= p.N) {
#pragma unroll
        for (int e = 0; e < 16; e += 4)
          *reinterpret_cast<float4*>(drow + ncol0 + c0 + e) = make_float4(o[e], o[e + 1], o[e + 2], o[e + 3]);
      } else {
#pragma unroll
        for (int e = 0; e < 16; ++e)
          if (ncol0 + c0 + e < p.N) drow[ncol0 + c0 + e] = o[e];
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(TMEM_COLS) : "memory");
}

// split-K: D = sum_z partial[z] (+bias)(ReLU), fixed order.  A block owns 32 consecutive output elements;
// its 8 warps each add every 8th partial (coalesced 128-byte reads), then warp 0 adds the 8 sums in order.
constexpr int kReduceElems = 32, kReduceLanes = 8;
__global__ void __launch_bounds__(kReduceElems * kReduceLanes)
xw_reduce_kernel(const float* __restrict__ part, int splits, int64_t M, int N, float* __restrict__ D,
                 int64_t ldd, const float* __restrict__ bias, int epi) {
  __shared__ float sums[kReduceLanes][kReduceElems];
  const int e = threadIdx.x % kReduceElems, zl = threadIdx.x / kReduceElems;
  const int64_t total = M * N;
  const int64_t idx = (int64_t)blockIdx.x * kReduceElems + e;
  float v = 0.f;
  if (idx < total)
    for (int z = zl; z < splits; z += kReduceLanes) v += part[(int64_t)z * total + idx];
  sums[zl][e] = v;
  __syncthreads();
  if (zl != 0 || idx >= total) return;
#pragma unroll
  for (int j = 1; j < kReduceLanes; ++j) v += sums[j][e];
  const int64_t r = idx / N;
  const int c = (int)(idx % N);
  if (epi & GCNB_EPI_BIAS) v += __ldg(bias + c);
  if (epi & GCNB_EPI_RELU) v = fmaxf(v, 0.f);
  D[r * ldd + c] = v;
}

}  // namespace xw
}  // namespace gcnb

using namespace gcnb;

namespace {

template <int BN, int MODE, bool BRAW>
int launch_main(const xw::Params& p, int64_t tiles, int splits, cudaStream_t stream) {
  // the > 48 KB shared-memory opt-in is a per-DEVICE attribute of the function: one bit per device
  // ordinal (a process may call gcnb_xw on cuda:1 after cuda:0).  Benign race: the attribute is idempotent.
  static std::atomic<uint64_t> configured{0};
  int dev = 0;
  GCNB_CHECK_CUDA(cudaGetDevice(&dev));
  const uint64_t bit = 1ull << (dev & 63);
  if (dev >= 64 || !(configured.load(std::memory_order_relaxed) & bit)) {
    GCNB_CHECK_CUDA(cudaFuncSetAttribute(xw::xw_kernel<BN, MODE, BRAW>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         xw::smem_bytes(BN, BRAW)));
    configured.fetch_or(bit, std::memory_order_relaxed);
  }
  xw::xw_kernel<BN, MODE, BRAW>
      <<<dim3((unsigned)tiles, (unsigned)splits), xw::THREADS, xw::smem_bytes(BN, BRAW), stream>>>(p);
  GCNB_CHECK_CUDA(cudaGetLastError());
  return GCNB_OK;
}

template <int MODE>
int launch_bn(int bn, bool braw, const xw::Params& p, int64_t tiles, int splits, cudaStream_t stream) {
  if (braw) {   // tall-K variant: bn <= 128
    switch (bn) {
      case 16: return launch_main<16, MODE, true>(p, tiles, splits, stream);
      case 32: return launch_main<32, MODE, true>(p, tiles, splits, stream);
      case 64: return launch_main<64, MODE, true>(p, tiles, splits, stream);
      default: return launch_main<128, MODE, true>(p, tiles, splits, stream);
    }
  }
  switch (bn) {
    case 16: return launch_main<16, MODE, false>(p, tiles, splits, stream);
    case 32: return launch_main<32, MODE, false>(p, tiles, splits, stream);
    case 64: return launch_main<64, MODE, false>(p, tiles, splits, stream);
    case 256: return launch_main<256, MODE, false>(p, tiles, splits, stream);
    default: return launch_main<128, MODE, false>(p, tiles, splits, stream);
  }
}

inline bool aligned16(const void* p) { return ((uintptr_t)p & 15) == 0; }

}  // namespace

namespace gcnb {
int xw_stream_run(const float* X, int64_t ldx, const float* W, int64_t ldw, int trans_w, float* D, int64_t ldd,
                  int64_t M, int N, int K, const float* bias, int epi, const uint32_t* mask_in, uint32_t* mask_out,
                  int64_t ldm, float drop_p, uint32_t seed, int precision, int n_sm, cudaStream_t stream);   // xw_stream.cu
int xw_tallk_stream_run(const float* X, int64_t ldx, const float* W, int64_t ldw, float* D, int64_t ldd, int64_t M,
                        int N, int64_t K, const float* bias, int epi, int precision, int n_sm, float* colsum,
                        cudaStream_t stream);                                                              // xw_tallk_stream.cu
// column sums of a row-major [K x N] matrix in two fixed-order passes (the fall-back of gcnb_xw_tn_colsum when
// the product kernel has no spare row for the ones column): slabs of kColsumRows rows -> part[slab][N] -> out[N].
// Every thread owns 4 columns of one row out of 8 at a time, four rows in flight; the second pass sums a column's
// slab values with 256 threads in a fixed order (strided partial sums, then a shared-memory tree).
constexpr int kColsumRows = 512;
__global__ void __launch_bounds__(256)
colsum_slab_kernel(const float* __restrict__ G, int64_t ldg, int64_t K, int N, float* __restrict__ part) {
  __shared__ float sh[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t r0 = (int64_t)blockIdx.x * kColsumRows, r1 = r0 + kColsumRows < K ? r0 + kColsumRows : K;
  for (int c0 = 0; c0 < N; c0 += 32) {
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    if (c0 + tx < N) {
      const float* g = G + c0 + tx;
      int64_t r = r0 + ty;
      for (; r + 24 < r1; r += 32) {
        s0 += g[r * ldg]; s1 += g[(r + 8) * ldg]; s2 += g[(r + 16) * ldg]; s3 += g[(r + 24) * ldg];
      }
      for (; r < r1; r += 8) s0 += g[r * ldg];
    }
    sh[ty][tx] = (s0 + s1) + (s2 + s3);
    __syncthreads();
    if (ty == 0 && c0 + tx < N) {
      float t = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) t += sh[j][tx];
      part[(int64_t)blockIdx.x * N + c0 + tx] = t;
    }
    __syncthreads();
  }
}
__global__ void __launch_bounds__(256)
colsum_final_kernel(const float* __restrict__ part, int slabs, int N, float* __restrict__ out) {
  __shared__ float sh[256];
  const int c = blockIdx.x;                      // one block per column
  float s = 0.f;
  for (int b = threadIdx.x; b < slabs; b += 256) s += part[(int64_t)b * N + c];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[c] = sh[0];
}
// elementwise tail of gcnb_xw_ex for the products that do not run on the streaming kernel: one thread per
// (row, 32-column word of the bit masks); same dropout hash (one word per 4 columns, one byte per element)
__global__ void __launch_bounds__(256)
xw_tail_kernel(float* __restrict__ D, int64_t ldd, const uint32_t* __restrict__ mask_in, uint32_t* __restrict__ mask_out,
               int64_t ldm, int64_t M, int N, int epi, float scale, uint32_t thresh, uint32_t seed) {
  const int words = (N + 31) / 32;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= M * words) return;
  const int64_t r = idx / words;
  const int w = (int)(idx % words);
  const int64_t n4 = (N + 3) / 4;
  const uint32_t min_word = (epi & GCNB_EPI_MASK) ? mask_in[r * ldm + w] : 0xffffffffu;
  uint32_t out_word = 0;
  for (int j = 0; j < 32; ++j) {
    const int c = w * 32 + j;
    if (c >= N) break;
    float x = D[r * ldd + c];
    if (epi & GCNB_EPI_DROPOUT) {
      const uint64_t g = (uint64_t)r * (uint64_t)n4 + (uint64_t)(c >> 2);
      uint32_t a = (uint32_t)g ^ seed, b = (uint32_t)(g >> 32) + 0x9e3779b9u * (seed | 1u);
      a *= 0x85ebca6bu; a ^= a >> 13; a += b; a *= 0xc2b2ae35u; a ^= a >> 16; a *= 0x27d4eb2fu; a ^= a >> 15;
      x = ((a >> (8 * (c & 3))) & 0xffu) >= thresh ? x * scale : 0.f;
    }
    if (epi & GCNB_EPI_MASK) x = ((min_word >> j) & 1u) ? x * scale : 0.f;
    if (x > 0.f) out_word |= 1u << j;
    D[r * ldd + c] = x;
  }
  if (mask_out) mask_out[r * ldm + w] = out_word;
}
}  // namespace gcnb

extern "C" int gcnb_xw_ex(const float* X, int64_t ldx, int trans_x, const float* W, int64_t ldw, int trans_w,
                          float* D, int64_t ldd, int64_t M, int N, int K, const float* bias, int epilogue,
                          const uint32_t* mask_in, uint32_t* mask_out, int64_t ldm, float drop_p, uint32_t seed,
                          int precision, void* stream_);

extern "C" int gcnb_xw(const float* X, int64_t ldx, int trans_x, const float* W, int64_t ldw, int trans_w,
                       float* D, int64_t ldd, int64_t M, int N, int K, const float* bias, int epilogue,
                       int precision, void* stream_) {
  GCNB_REQUIRE((epilogue & ~(GCNB_EPI_BIAS | GCNB_EPI_RELU)) == 0, "unknown epilogue flags %d", epilogue);
  return gcnb_xw_ex(X, ldx, trans_x, W, ldw, trans_w, D, ldd, M, N, K, bias, epilogue, nullptr, nullptr, 0, 0.f, 0,
                    precision, stream_);
}

extern "C" int gcnb_xw_ex(const float* X, int64_t ldx, int trans_x, const float* W, int64_t ldw, int trans_w,
                          float* D, int64_t ldd, int64_t M, int N, int K, const float* bias, int epilogue_ex,
                          const uint32_t* mask_in, uint32_t* mask_out, int64_t ldm, float drop_p, uint32_t seed,
                          int precision, void* stream_) {
  GCNB_REQUIRE(M >= 0 && N >= 0 && K >= 0, "negative size (M=%lld N=%d K=%d)", (long long)M, N, K);
  GCNB_REQUIRE(precision >= 0 && precision <= 2, "precision %d not in 0..2", precision);
  GCNB_REQUIRE((epilogue_ex & ~(GCNB_EPI_BIAS | GCNB_EPI_RELU | GCNB_EPI_DROPOUT | GCNB_EPI_MASK)) == 0,
               "unknown epilogue flags %d", epilogue_ex);
  GCNB_REQUIRE(!(epilogue_ex & GCNB_EPI_MASK) || mask_in != nullptr, "GCNB_EPI_MASK without mask_in");
  GCNB_REQUIRE((mask_in == nullptr && mask_out == nullptr) || ldm >= (N + 31) / 32, "ldm smaller than ceil(N/32) words");
  GCNB_REQUIRE(drop_p >= 0.f && drop_p < 1.f, "dropout probability %f not in [0, 1)", (double)drop_p);
  const int epilogue = epilogue_ex & (GCNB_EPI_BIAS | GCNB_EPI_RELU);
  const int tail = (epilogue_ex & (GCNB_EPI_DROPOUT | GCNB_EPI_MASK)) | (mask_out ? 0x100 : 0);
  GCNB_REQUIRE(!(epilogue & GCNB_EPI_BIAS) || bias != nullptr, "GCNB_EPI_BIAS without a bias pointer");
  if (M == 0 || N == 0) return GCNB_OK;
  GCNB_REQUIRE(D != nullptr && ldd >= N, "D is null or ldd < N");
  GCNB_REQUIRE(K == 0 || (X != nullptr && W != nullptr), "X or W is null");
  GCNB_REQUIRE(K == 0 || (ldx >= (trans_x ? M : (int64_t)K) && ldw >= (trans_w ? K : N)),
               "leading dimension smaller than the stored row length");
  GCNB_REQUIRE(M < ((int64_t)1 << 37), "M too large");
  int rc = require_gpu();
  if (rc != GCNB_OK) return rc;
  const int dev = pointer_device(D);
  GCNB_REQUIRE(dev >= 0, "D is not a device pointer");
  DeviceGuard guard(dev);
  GCNB_CHECK_CUDA(guard.status);
  cudaStream_t stream = (cudaStream_t)stream_;
  // elementwise tail (dropout / mask) as its own pass, for the products the streaming kernel does not take
  auto run_tail = [&]() -> int {
    if (!tail) return GCNB_OK;
    const int64_t total = M * ((N + 31) / 32);
    int t8 = (int)(drop_p * 256.f + 0.5f);
    t8 = t8 < 0 ? 0 : (t8 > 255 ? 255 : t8);
    xw_tail_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(D, ldd, mask_in, mask_out, ldm, M, N, tail & 0xff,
                                                                          256.f / (256.f - (float)t8), (uint32_t)t8, seed);
    GCNB_CHECK_CUDA(cudaGetLastError());
    return GCNB_OK;
  };

  if (K > 0 && !trans_x) {
    const int r = xw_stream_run(X, ldx, W, ldw, trans_w, D, ldd, M, N, K, bias, epilogue_ex, mask_in, mask_out, ldm,
                                drop_p, seed, precision, sm_count(dev), stream);
    if (r != 1) return r;   // ran (or failed) there; 1 = not that kernel's shape
  }

  if (K > 0 && trans_x && !trans_w) {   // dW = X^T·G with both operands row-major over the vertices
    const int r = xw_tallk_stream_run(X, ldx, W, ldw, D, ldd, M, N, K, bias, epilogue, precision, sm_count(dev), nullptr, stream);
    if (r == GCNB_OK) return run_tail();
    if (r != 1) return r;
  }

  if (K == 0) {   // empty sum: D = epilogue(0)
    const int64_t total = M * N;
    xw::xw_reduce_kernel<<<(unsigned)((total + xw::kReduceElems - 1) / xw::kReduceElems), xw::kReduceElems * xw::kReduceLanes, 0, stream>>>(
        nullptr, 0, M, N, D, ldd, bias, epilogue);
    GCNB_CHECK_CUDA(cudaGetLastError());
    return run_tail();
  }

  // N > 128: a 256-wide tile would convert X once instead of once per 128-wide n-tile, but its 131 KB of
  // shared memory leave one 4-warp CTA per SM: measured slower (products epoch 23.3 vs 21.9 ms,
  // profiles/r01_ae_*), so it stays an experiment switch
  static const int wide = getenv("GCNB_XW_BN256") ? atoi(getenv("GCNB_XW_BN256")) : 0;
  const int64_t mtiles = (M + xw::BM - 1) / xw::BM;
  // tall K, both operands vertex-sized (dW = X^T·dS): convert the second operand in the kernel instead
  // of writing and re-reading an image of it.  Read per call so that a test can compare the two paths.
  const char* tallk_env = getenv("GCNB_XW_TALLK");
  const bool braw = (tallk_env ? atoi(tallk_env) != 0 : true) && !trans_w && mtiles <= 4 && K >= 4096;
  const char* tallk_bn_env = getenv("GCNB_XW_TALLK_BN");   // experiment: 64 = two CTAs per SM, X converted per n-tile
  const int wide_bn = braw ? ((tallk_bn_env && atoi(tallk_bn_env) == 64) ? 64 : 128) : ((N > 128 && wide) ? 256 : 128);
  const int bn = N <= 16 ? 16 : (N <= 32 ? 32 : (N <= 64 ? 64 : wide_bn));
  const int ntiles = (N + bn - 1) / bn;
  const int nkb = (K + xw::BK - 1) / xw::BK;
  const int64_t tiles = mtiles * ntiles;
  GCNB_REQUIRE(tiles < 0x7fffffffLL, "too many tiles");
  const int n_sm = sm_count(dev);
  // CTAs the split-K grid aims for: two per SM, or what the tall-K variant's shared memory leaves room for
  const int ctas_per_sm = braw ? std::max(1, std::min(2, (227 * 1024) / (xw::smem_bytes(bn, true) + 1024))) : 2;
  const int64_t want = (int64_t)ctas_per_sm * n_sm;
  int splits = 1;
  if (tiles < want && nkb > 1) splits = (int)std::min<int64_t>(nkb, (want + tiles - 1) / tiles);
  if (braw || (K >= 4096 && mtiles <= 8)) {   // tall K, small output: partials are cheap
    // The TMEM accumulator does not round to nearest: the error of one split grows with the number of
    // MMAs chained into it (measured: K = 33 333 in 25 splits, 1.05e-5 of max|D|; K = 2.4 M in 74 vs 148
    // splits differ by 1.3e-4).  Bound the chain; the partial sums are added in fp32 round-to-nearest.
    // The partials stay under 1 GB.
    const char* chain_env = getenv("GCNB_XW_MAX_CHAIN");
    const int64_t max_chain = std::max(1, chain_env ? atoi(chain_env) : 64);   // k-blocks of 16 per split
    const int64_t max_splits = std::max<int64_t>(1, std::min<int64_t>(65535, ((int64_t)1 << 28) / std::max<int64_t>(1, M * N)));
    const int64_t by_chain = std::min<int64_t>(max_splits, (nkb + max_chain - 1) / max_chain);
    splits = (int)std::max<int64_t>(splits, by_chain);
  }
  const int kb_per_split = (nkb + splits - 1) / splits;
  splits = (nkb + kb_per_split - 1) / kb_per_split;
  GCNB_REQUIRE(splits <= 65535, "too many k splits");

  uint8_t* img = nullptr;
  float* partial = nullptr;
  const size_t img_bytes = (size_t)ntiles * nkb * 2 * xw::b_part_bytes(bn);
  if (!braw) {
    GCNB_CHECK_CUDA(cudaMallocAsync(&img, img_bytes, stream));
    if (precision != 0) GCNB_CHECK_CUDA(cudaMemsetAsync(img, 0, img_bytes, stream));  // unused lo parts stay defined
    const int64_t total = (int64_t)ntiles * bn * nkb * xw::BK;
    const unsigned blocks = (unsigned)((total + 255) / 256);
    if (precision == 0) xw::prep_w_kernel<0><<<blocks, 256, 0, stream>>>(W, ldw, trans_w, N, K, bn, nkb, ntiles, img);
    else if (precision == 1) xw::prep_w_kernel<1><<<blocks, 256, 0, stream>>>(W, ldw, trans_w, N, K, bn, nkb, ntiles, img);
    else xw::prep_w_kernel<2><<<blocks, 256, 0, stream>>>(W, ldw, trans_w, N, K, bn, nkb, ntiles, img);
    GCNB_CHECK_CUDA(cudaGetLastError());
  }
  xw::Params p;
  p.X = X; p.ldx = ldx; p.trans_x = trans_x ? 1 : 0; p.w_img = img;
  p.W = W; p.ldw = ldw; p.w_vec = aligned16(W) && (ldw % 4 == 0);
  p.M = M; p.N = N; p.K = K; p.nkb = nkb; p.ntiles = ntiles; p.kb_per_split = kb_per_split;
  p.x_vec = aligned16(X) && (ldx % 4 == 0);
  if (splits > 1) {
    GCNB_CHECK_CUDA(cudaMallocAsync(&partial, (size_t)splits * M * N * sizeof(float), stream));
    p.D = partial; p.ldd = N; p.bias = nullptr; p.epi = 0;
    p.d_vec = (N % 4 == 0);
  } else {
    p.D = D; p.ldd = ldd; p.bias = bias; p.epi = epilogue;
    p.d_vec = aligned16(D) && (ldd % 4 == 0);
  }
  if (precision == 0) rc = launch_bn<0>(bn, braw, p, tiles, splits, stream);
  else if (precision == 1) rc = launch_bn<1>(bn, braw, p, tiles, splits, stream);
  else rc = launch_bn<2>(bn, braw, p, tiles, splits, stream);
  if (rc == GCNB_OK && splits > 1) {
    const int64_t total = M * N;
    xw::xw_reduce_kernel<<<(unsigned)((total + xw::kReduceElems - 1) / xw::kReduceElems), xw::kReduceElems * xw::kReduceLanes, 0, stream>>>(
        partial, splits, M, N, D, ldd, bias, epilogue);
    if (cudaGetLastError() != cudaSuccess) rc = set_error(GCNB_ERR_CUDA, "xw_reduce_kernel launch failed");
  }
  if (img) cudaFreeAsync(img, stream);
  if (partial) cudaFreeAsync(partial, stream);
  if (rc == GCNB_OK) rc = run_tail();
  return rc;
}

extern "C" int gcnb_xw_tn_colsum(const float* X, int64_t ldx, const float* G, int64_t ldg, float* D, int64_t ldd,
                                 float* colsum, int64_t M, int N, int K, int precision, void* stream_) {
  GCNB_REQUIRE(M >= 0 && N >= 0 && K >= 0, "negative size (M=%lld N=%d K=%d)", (long long)M, N, K);
  GCNB_REQUIRE(precision >= 0 && precision <= 2, "precision %d not in 0..2", precision);
  GCNB_REQUIRE(N == 0 || colsum != nullptr, "colsum is null");
  if (N == 0) return GCNB_OK;
  int rc = require_gpu();
  if (rc != GCNB_OK) return rc;
  const int dev = pointer_device(colsum);
  GCNB_REQUIRE(dev >= 0, "colsum is not a device pointer");
  DeviceGuard guard(dev);
  GCNB_CHECK_CUDA(guard.status);
  cudaStream_t stream = (cudaStream_t)stream_;
  if (M > 0 && K > 0) {
    GCNB_REQUIRE(D != nullptr && ldd >= N && X != nullptr && G != nullptr && ldx >= M && ldg >= N,
                 "null operand or leading dimension smaller than the stored row length");
    const int r = xw_tallk_stream_run(X, ldx, G, ldg, D, ldd, M, N, K, nullptr, 0, precision, sm_count(dev), colsum, stream);
    if (r != 1) return r;   // both results came out of the one pass (or it failed)
  }
  if (M > 0) {
    rc = gcnb_xw_ex(X, ldx, 1, G, ldg, 0, D, ldd, M, N, K, nullptr, 0, nullptr, nullptr, 0, 0.f, 0, precision, stream_);
    if (rc != GCNB_OK) return rc;
  }
  if (K == 0) {
    GCNB_CHECK_CUDA(cudaMemsetAsync(colsum, 0, (size_t)N * sizeof(float), stream));
    return GCNB_OK;
  }
  GCNB_REQUIRE(G != nullptr && ldg >= N, "G is null or ldg < N");
  const int slabs = (int)(((int64_t)K + kColsumRows - 1) / kColsumRows);
  float* part = nullptr;
  GCNB_CHECK_CUDA(cudaMallocAsync(&part, (size_t)slabs * N * sizeof(float), stream));
  colsum_slab_kernel<<<slabs, 256, 0, stream>>>(G, ldg, K, N, part);
  rc = cudaGetLastError() == cudaSuccess ? GCNB_OK : set_error(GCNB_ERR_CUDA, "colsum_slab_kernel launch failed");
  if (rc == GCNB_OK) {
    colsum_final_kernel<<<N, 256, 0, stream>>>(part, slabs, N, colsum);
    if (cudaGetLastError() != cudaSuccess) rc = set_error(GCNB_ERR_CUDA, "colsum_final_kernel launch failed");
  }
  cudaFreeAsync(part, stream);
  return rc;
}

// ---- column sums without any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA-graph capture, tight serving loops) -------------------
// The slab partials of colsum_slab_kernel live in a file-scope device array instead of a pool allocation.  Every
// device that loads the library holds these 8 MiB until exit.  The array name below is a HOST-side stand-in in host
// code: its device address is looked up once per device with cudaGetSymbolAddress and that pointer is what the
// kernels receive.  All calls share the array: two calls that overlap on one GPU (different streams) race.
// Nothing needs clearing: the first pass writes every part[slab][0..N) the second pass reads.
namespace {
constexpr int64_t kNoAllocScratchFloats = (int64_t)2 << 20;              // 8 MiB
constexpr int kNoAllocMaxDevices = 64;
static __device__ __align__(256) float g_colsum_scratch[kNoAllocScratchFloats];
std::atomic<float*> g_colsum_scratch_addr[kNoAllocMaxDevices];

// device address of the scratch array on `dev` (which the caller has made current); nullptr + error on failure
int colsum_scratch_for(int dev, float** out) {
  GCNB_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, "device %d beyond the %d devices the no-alloc table holds", dev,
               kNoAllocMaxDevices);
  float* p = g_colsum_scratch_addr[dev].load(std::memory_order_acquire);
  if (p == nullptr) {
    void* sym = nullptr;
    GCNB_CHECK_CUDA(cudaGetSymbolAddress(&sym, g_colsum_scratch));
    GCNB_REQUIRE(sym != nullptr, "cudaGetSymbolAddress returned a null scratch address on device %d", dev);
    p = static_cast<float*>(sym);
    g_colsum_scratch_addr[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return GCNB_OK;
}
}  // namespace

extern "C" int64_t gcnb_colsum_noalloc_max_floats(void) { return kNoAllocScratchFloats; }

extern "C" int gcnb_colsum_noalloc(const float* G, int64_t ldg, int64_t K, int N, float* colsum, void* stream_) {
  GCNB_REQUIRE(K >= 0 && N >= 0, "negative size (K=%lld N=%d)", (long long)K, N);
  GCNB_REQUIRE(N == 0 || colsum != nullptr, "colsum is null");
  if (N == 0) return GCNB_OK;
  const int64_t slabs = (K + kColsumRows - 1) / kColsumRows;
  GCNB_REQUIRE(slabs * (int64_t)N <= kNoAllocScratchFloats,
               "gcnb_colsum_noalloc: ceil(K/%d)*N = %lld partial sums exceed the %lld-float static scratch "
               "(K=%lld N=%d); use gcnb_xw_tn_colsum",
               kColsumRows, (long long)(slabs * (int64_t)N), (long long)kNoAllocScratchFloats, (long long)K, N);
  int rc = require_gpu();
  if (rc != GCNB_OK) return rc;
  const int dev = pointer_device(colsum);
  GCNB_REQUIRE(dev >= 0, "colsum is not a device pointer");
  DeviceGuard guard(dev);
  GCNB_CHECK_CUDA(guard.status);
  cudaStream_t stream = (cudaStream_t)stream_;
  if (K == 0) {
    GCNB_CHECK_CUDA(cudaMemsetAsync(colsum, 0, (size_t)N * sizeof(float), stream));
    return GCNB_OK;
  }
  GCNB_REQUIRE(G != nullptr && ldg >= N, "G is null or ldg < N");
  float* part = nullptr;
  rc = colsum_scratch_for(dev, &part);
  if (rc != GCNB_OK) return rc;
  colsum_slab_kernel<<<(unsigned)slabs, 256, 0, stream>>>(G, ldg, K, N, part);
  if (cudaGetLastError() != cudaSuccess) return set_error(GCNB_ERR_CUDA, "colsum_slab_kernel launch failed");
  colsum_final_kernel<<<N, 256, 0, stream>>>(part, (int)slabs, N, colsum);
  if (cudaGetLastError() != cudaSuccess) return set_error(GCNB_ERR_CUDA, "colsum_final_kernel launch failed");
  return GCNB_OK;
}
