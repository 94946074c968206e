// Weight-gradient product of the GCN layer, dW = X^T · G, as a persistent warp-specialised tcgen05 kernel:
// D[M x N] = sum over v of A[v, 0..M)^T · B[v, 0..N) with v = vertices (millions), M <= 128, N <= 256 —
// what autograd derives from torch.mm(input, self.weight) (pygcn/gcn6.py:131, :191) for the weight.
//
// Both operands are vertex-sized row-major matrices [K x M] and [K x N]: the reduction runs over their ROWS.
// The first kernel's tall-K variant (xw_tcgen05.cu, BRAW) transposes them into K-major tiles in software
// (4-byte transposing shared-memory stores, one 4-warp CTA) and was instruction-bound: 3.5 + 1.7 ms of the
// 16.6 ms products-shaped epoch for 6.5 GB of operands (profiles/r02_f_launches_epoch_products.csv).
// Blackwell's tensor core takes MN-major operands for tf32 (layout SWIZZLE_128B_BASE32B: rows = k, 128-byte rows
// of 32 consecutive m), so no transposition is needed, only a re-layout of whole 16-byte chunks.
// The first version of this kernel let TMA tensor-map boxes (16 vertices x 32 columns, SWIZZLE_128B_ATOM_32B)
// write that layout directly; this version loads whole rows with 1-D bulk copies (fewer, larger requests; both
// stream at the HBM rate when >= 64 KB are in flight, profiles/r02_s_tma_stream_probe.jsonl — what made the first
// version slow was its 2 516 partial matrices, DESIGN.md "Round 2: the dense products"):
//   warp 0      producer: per stage of 16 vertices ONE 1-D bulk copy per operand (16 x M x 4 contiguous bytes when
//               the matrix is dense in memory, else one per row) into a ring of raw stages
//   warps 4-11  converters: raw rows -> hi = trunc13(x) and lo = tf32(x - hi) tiles of BOTH operands, written
//               as 16-byte chunks at the swizzled position of the MN-major layout (ring of 2 converted stages)
//   warp 1      MMA issuer: per stage 2 k-steps x (hi·hi + hi·lo + lo·hi), M = 128 x N, accumulator in TMEM;
//               the TMEM accumulator does not round to nearest, so a chain is bounded at
//               `chain` stages (64 = 1024 vertices) and the partial is handed to the epilogue; two buffers
//   warps 12-19 epilogue: tcgen05.ld -> partial matrix in global memory (fp32, round-to-nearest sums later)
//   The partials of all CTAs are added in a fixed order by xt_reduce_kernel (also transposes the result
//   when the caller's M exceeded 128 and the product was computed as (G^T · X)^T).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include <algorithm>
#include <atomic>

#include "common.h"

namespace gcnb {
namespace xt {

constexpr int KV = 16;                 // vertices (k) per stage
constexpr int BLK = KV * 128;          // bytes of one 32-column block of a converted stage (KV rows x 128 B)
constexpr int A_BLOCKS = 4, B_BLOCKS = 8;
constexpr int PART = (A_BLOCKS + B_BLOCKS) * BLK;    // hi (or lo) tiles of one stage: 24 KB
constexpr int CONV_STAGE = 2 * PART;                 // hi | lo
constexpr int CS = 2;                  // converted stages in flight
constexpr int RAW_MAX = 8;             // raw stages in flight (barrier slots)
constexpr int NC = 8, NE = 8;          // converter / epilogue warps
constexpr int THREADS = (4 + NC + NE) * 32;
constexpr int SMEM_MAX = 232448;

struct Params {
  const float* A; int64_t lda;   // [K x M] row-major, the operand whose columns become the rows of D
  const float* B; int64_t ldb;   // [K x N] row-major
  float* part;               // [grid][rows][N] one partial matrix per CTA (rows = M, +1 with the ones column)
  int64_t K;                 // vertices
  int M, N;                  // used rows / columns (M <= 128, N <= 256, both multiples of 4)
  int M1;                    // rows of a partial matrix: M, or ones_col + 1
  int mb, nb;                // 32-column blocks of A / B in use
  int64_t v_per_cta;         // vertices per CTA (multiple of KV)
  int chain;                 // stages per accumulation chain in TMEM
  int ones_col;              // -1, or the column of A (>= M, < 128) the converters fill with 1.0: row `ones_col`
                             // of the product is then the column sums of B (the bias gradient next to dW)
  int raws;                  // raw stages in flight
  int raw_stage;             // bytes of one raw stage: KV x (M + N) floats, rounded up to 128
};

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
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
__device__ __forceinline__ void mbar_wait_sleep(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) break;
    __nanosleep(256);
  }
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// MN-major tf32 operand, layout SWIZZLE_128B_BASE32B (type 1): rows = k, 128 bytes (32 consecutive m) per row, the
// 32-byte units of a row XOR-ed with (row & 3); 4-row k-atoms 512 bytes apart; `lbo` = bytes between consecutive
// 32-element blocks in M/N.  Pinned on hardware (profiles/r02_j_xt_variants.jsonl): with the plain SWIZZLE_128B
// layout the tensor core returns zeros for MN-major tf32.
__device__ __forceinline__ uint64_t make_desc_mn(uint32_t saddr, uint32_t lbo) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3fff);
  d |= (uint64_t)((lbo >> 4) & 0x3fff) << 16;
  d |= (uint64_t)(512 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)1 << 61;
  return d;
}
// byte offset of the 16-byte chunk (k-row v, chunk c16 of the 128-byte row) inside a block
__device__ __forceinline__ uint32_t mn_off(int v, int c16) {
  return (uint32_t)(v * 128 + ((((c16 >> 1) ^ (v & 3)) << 5) | ((c16 & 1) << 4)));
}
// D fp32, A/B TF32, both MN-major (bits 15, 16), M = 128, N = n
__device__ __forceinline__ uint32_t make_idesc(int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}
__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool elect_one() {   // one lane of a converged warp
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ uint32_t to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}

template <int MODE>   // 0: 3xTF32, 1: single-pass TF32
__global__ void __launch_bounds__(THREADS, 1)
xt_kernel(const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_addr(smem_raw) & 1023u)) & 1023u);
  uint8_t* conv = smem;                        // CS stages: { hi: A blocks 0..3 | B blocks 0..7 } { lo: same }
  uint8_t* raw = conv + CS * CONV_STAGE;       // raws stages: A rows [KV][M] then B rows [KV][N], row-major
  uint64_t* bars = reinterpret_cast<uint64_t*>(raw + p.raws * p.raw_stage);
  // bars: FULL[RAW_MAX] EMPTY[RAW_MAX] CONV[CS] CEMPTY[CS] TFULL[2] TEMPTY[2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * RAW_MAX + 2 * CS + 4);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t bar0 = smem_addr(bars);
  auto FULL = [&](int s) { return bar0 + 8u * (uint32_t)s; };
  auto EMPTY = [&](int s) { return bar0 + 8u * (uint32_t)(RAW_MAX + s); };
  auto CONV = [&](int s) { return bar0 + 8u * (uint32_t)(2 * RAW_MAX + s); };
  auto CEMPTY = [&](int s) { return bar0 + 8u * (uint32_t)(2 * RAW_MAX + CS + s); };
  auto TFULL = [&](int b) { return bar0 + 8u * (uint32_t)(2 * RAW_MAX + 2 * CS + b); };
  auto TEMPTY = [&](int b) { return bar0 + 8u * (uint32_t)(2 * RAW_MAX + 2 * CS + 2 + b); };
  const int n_umma = p.nb * 32;                // UMMA N (multiple of 32, <= 256)
  constexpr int TMEM_COLS = 512;               // two accumulator buffers of up to 256 columns
  const int RAWS = p.raws;

  if (tid == 0) {
    for (int s = 0; s < RAW_MAX; ++s) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(FULL(s)));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(EMPTY(s)), "r"(NC * 32));
    }
    for (int s = 0; s < CS; ++s) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(CONV(s)), "r"(NC * 32));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(CEMPTY(s)));
    }
    for (int b = 0; b < 2; ++b) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(TFULL(b)));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(TEMPTY(b)), "r"(NE * 32));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // the converted tiles are zeroed once: the converters never write the columns past M / N of the last block
  // nor the blocks past mb / nb (the MMA always runs at M = 128), so those stay zero for the whole kernel
  for (int i = tid; i < CS * CONV_STAGE / 16; i += THREADS) reinterpret_cast<uint4*>(conv)[i] = make_uint4(0, 0, 0, 0);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_addr(tmem_slot)), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  const int64_t v0 = (int64_t)blockIdx.x * p.v_per_cta;
  const int64_t v1 = v0 + p.v_per_cta < p.K ? v0 + p.v_per_cta : p.K;
  const int nst = v1 > v0 ? (int)((v1 - v0 + KV - 1) / KV) : 0;    // stages of this CTA
  const int ngroups = (nst + p.chain - 1) / p.chain;                 // partials it produces

  if (warp == 0) {
    // ---------------- producer: 1-D bulk copies of whole rows (one per operand and stage when the rows are
    // contiguous, else one per row: small copies cost ~300 cycles each, tools/tma_stream_probe.cu) ----------------
    if (lane == 0) {
      const bool a_contig = p.lda == p.M, b_contig = p.ldb == p.N;
      int s = 0;               // ring position / phase as counters (no integer division per stage)
      uint32_t ph = 0;
      for (int it = 0; it < nst; ++it, s = (s + 1 == RAWS ? 0 : s + 1), ph ^= (s == 0)) {
        mbar_wait(EMPTY(s), ph ^ 1);
        const int64_t v = v0 + (int64_t)it * KV;
        const int rows = (int)(v1 - v < KV ? v1 - v : KV);
        const uint32_t full = FULL(s);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(full), "r"((uint32_t)(rows * (p.M + p.N) * 4)) : "memory");
        const uint32_t dst_a = smem_addr(raw + s * p.raw_stage), dst_b = dst_a + (uint32_t)(KV * p.M * 4);
        if (a_contig) {
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                       ::"r"(dst_a), "l"(p.A + v * p.lda), "r"((uint32_t)(rows * p.M * 4)), "r"(full) : "memory");
        } else {
          for (int r = 0; r < rows; ++r)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(dst_a + (uint32_t)(r * p.M * 4)), "l"(p.A + (v + r) * p.lda), "r"((uint32_t)(p.M * 4)), "r"(full) : "memory");
        }
        if (b_contig) {
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                       ::"r"(dst_b), "l"(p.B + v * p.ldb), "r"((uint32_t)(rows * p.N * 4)), "r"(full) : "memory");
        } else {
          for (int r = 0; r < rows; ++r)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(dst_b + (uint32_t)(r * p.N * 4)), "l"(p.B + (v + r) * p.ldb), "r"((uint32_t)(p.N * 4)), "r"(full) : "memory");
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer: the whole warp walks the loop (uniform control flow, descriptors in uniform
    // registers), one elected lane issues — under `if (lane == 0)` every tcgen05.mma is wrapped in a uniformising
    // loop of ~17 instructions (xw_stream.cu) ----------------
    {
      const uint32_t idesc = make_idesc(n_umma);
      int grp = 0, pos = 0;
      for (int it = 0; it < nst; ++it) {
        const uint32_t buf = grp & 1;
        if (pos == 0) mbar_wait(TEMPTY(buf), ((grp >> 1) & 1) ^ 1);
        const uint32_t tmem_d = tmem_base + buf * 256;
        const int cs = it % CS;
        mbar_wait(CONV(cs), (it / CS) & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t a_hi = smem_addr(conv + cs * CONV_STAGE), b_hi = a_hi + A_BLOCKS * BLK;
        const uint32_t a_lo = a_hi + PART, b_lo = a_lo + A_BLOCKS * BLK;
        __syncwarp();
        if (elect_one()) {
#pragma unroll
          for (int ks = 0; ks < KV / 8; ++ks) {        // 8 vertices per MMA: two 4-row k-atoms of every block
            const uint32_t o = ks * 1024;
            const uint32_t acc = (pos > 0 || ks > 0) ? 1u : 0u;
            mma_tf32(tmem_d, make_desc_mn(a_hi + o, BLK), make_desc_mn(b_hi + o, BLK), idesc, acc);
            if constexpr (MODE == 0) {
              mma_tf32(tmem_d, make_desc_mn(a_hi + o, BLK), make_desc_mn(b_lo + o, BLK), idesc, 1u);
              mma_tf32(tmem_d, make_desc_mn(a_lo + o, BLK), make_desc_mn(b_hi + o, BLK), idesc, 1u);
            }
          }
          commit(CEMPTY(cs));
          if (pos == p.chain - 1 || it == nst - 1) commit(TFULL(buf));
        }
        __syncwarp();
        if (++pos == p.chain) { pos = 0; ++grp; }
      }
    }
  } else if (warp >= 4 && warp < 4 + NC) {
    // ---------------- converters: raw rows -> MN-major hi / lo tiles (both operands) ----------------
    const int ct = tid - 128;                        // 0 .. NC*32-1
    const int mpc = p.M >> 2, npc = p.N >> 2;        // 16-byte chunks per row
    const uint32_t inv_m = (65536u + (uint32_t)mpc - 1) / (uint32_t)mpc, inv_n = (65536u + (uint32_t)npc - 1) / (uint32_t)npc;
    int s = 0;
    uint32_t ph = 0;
    for (int it = 0; it < nst; ++it) {
      const int cs = it % CS;
      mbar_wait(FULL(s), ph);
      mbar_wait(CEMPTY(cs), ((it / CS) & 1) ^ 1);
      const int64_t v = v0 + (int64_t)it * KV;
      const int rows = (int)(v1 - v < KV ? v1 - v : KV);
      const uint8_t* ra = raw + s * p.raw_stage;
      const uint8_t* rb = ra + KV * p.M * 4;
      uint8_t* hi = conv + cs * CONV_STAGE;
      uint8_t* lo = hi + PART;
      auto convert = [&](const uint8_t* src, int cols4, uint32_t inv, uint8_t* dhi, uint8_t* dlo) {
        for (int i = ct; i < KV * cols4; i += NC * 32) {
          const int vr = (int)(((uint32_t)i * inv) >> 16);          // i / cols4 (exact for i < 1024, cols4 <= 64)
          const int c = i - vr * cols4;
          uint4 x = make_uint4(0, 0, 0, 0);
          if (vr < rows) x = *reinterpret_cast<const uint4*>(src + (size_t)i * 16);   // rows past the end of K are zero
          const uint4 h = make_uint4(x.x & 0xffffe000u, x.y & 0xffffe000u, x.z & 0xffffe000u, x.w & 0xffffe000u);
          const uint32_t off = (uint32_t)((c >> 3) * BLK) + mn_off(vr, c & 7);
          *reinterpret_cast<uint4*>(dhi + off) = h;
          if constexpr (MODE == 0) {
            uint4 l;
            l.x = to_tf32(__uint_as_float(x.x) - __uint_as_float(h.x));
            l.y = to_tf32(__uint_as_float(x.y) - __uint_as_float(h.y));
            l.z = to_tf32(__uint_as_float(x.z) - __uint_as_float(h.z));
            l.w = to_tf32(__uint_as_float(x.w) - __uint_as_float(h.w));
            *reinterpret_cast<uint4*>(dlo + off) = l;
          }
        }
      };
      convert(ra, mpc, inv_m, hi, lo);
      convert(rb, npc, inv_n, hi + A_BLOCKS * BLK, lo + A_BLOCKS * BLK);
      if (p.ones_col >= 0 && ct < KV) {            // A[v, ones_col] = 1 for the vertices that exist (its low part stays 0)
        const int c = p.ones_col >> 2;
        *reinterpret_cast<float*>(hi + (c >> 3) * BLK + mn_off(ct, c & 7) + (p.ones_col & 3) * 4) = ct < rows ? 1.f : 0.f;
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic stores -> tensor core's proxy
      mbar_arrive(CONV(cs));
      mbar_arrive(EMPTY(s));                          // this thread's reads of the raw stage are done
      if (++s == RAWS) { s = 0; ph ^= 1; }
    }
  } else if (warp >= 4 + NC) {
    // ---------------- epilogue: accumulator -> partial matrix ----------------
    const int q = warp & 3, eg = (warp - 4 - NC) >> 2;
    const int row = q * 32 + lane;                   // row of D = column of A
    for (int grp = 0; grp < ngroups; ++grp) {
      const uint32_t buf = grp & 1;
      mbar_wait_sleep(TFULL(buf), (grp >> 1) & 1);   // a chain takes tens of microseconds: do not spin on the issue slots
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t taddr0 = tmem_base + buf * 256 + ((uint32_t)(q * 32) << 16);
      // one partial matrix per CTA, accumulated chain by chain by the thread that owns the row (L2-resident,
      // 100 KB per CTA): round-to-nearest fp32 sums in a fixed order, and the final reduction reads 148 partials
      float* dst = p.part + (int64_t)blockIdx.x * p.M1 * p.N + (int64_t)row * p.N;
#pragma unroll 1
      for (int c0 = eg * 32; c0 < n_umma; c0 += (NE / 4) * 32) {
        uint32_t r[32];
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
              "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
            : "r"(taddr0 + (uint32_t)c0) : "memory");
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
            : "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
              "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
            : "r"(taddr0 + (uint32_t)(c0 + 16)) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (row < p.M1) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            if (c0 + j < p.N) {                      // N is a multiple of 4
              float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
              if (grp > 0) o = *reinterpret_cast<const float4*>(dst + c0 + j);
              o.x += __uint_as_float(r[j]); o.y += __uint_as_float(r[j + 1]);
              o.z += __uint_as_float(r[j + 2]); o.w += __uint_as_float(r[j + 3]);
              *reinterpret_cast<float4*>(dst + c0 + j) = o;
            }
          }
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      mbar_arrive(TEMPTY(buf));
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
}

// D (+bias)(ReLU) = sum of the CTAs' partial matrices in a fixed order; `transposed`: the partials are [N x M] = D^T.
// Row M of the partials (ones column, direct route only) goes to colsum[0..N).  One thread per 4 consecutive
// elements of a row (Np is a multiple of 4), 8 independent loads in flight.
__global__ void __launch_bounds__(128)
xt_reduce_kernel(const float* __restrict__ part, int n_valid, int M1, int Mp, int Np, int transposed,
                 float* __restrict__ D, int64_t ldd, int M, int N, const float* __restrict__ bias, int epi, int ones_row,
                 float* __restrict__ colsum) {
  const int idx4 = blockIdx.x * blockDim.x + threadIdx.x;    // group of 4 elements of the partial matrix (M1 x Np)
  const int total = M1 * Np;
  if (idx4 * 4 >= total) return;
  const float4* src = reinterpret_cast<const float4*>(part) + idx4;
  const int64_t stride4 = (int64_t)total / 4;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
  for (int c = 0; c < n_valid; ++c) {
    const float4 v = __ldcs(src + (int64_t)c * stride4);
    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
  }
  const int pr = (idx4 * 4) / Np, pc = (idx4 * 4) % Np;
  const float o[4] = {s.x, s.y, s.z, s.w};
  if (pr >= Mp) {
    if (pr == ones_row && colsum) {
#pragma unroll
      for (int e = 0; e < 4; ++e) colsum[pc + e] = o[e];
    }
    return;
  }
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const int r = transposed ? pc + e : pr, c = transposed ? pr : pc + e;
    if (r >= M || c >= N) continue;
    float v = o[e];
    if (epi & GCNB_EPI_BIAS) v += __ldg(bias + c);
    if (epi & GCNB_EPI_RELU) v = fmaxf(v, 0.f);
    D[(int64_t)r * ldd + c] = v;
  }
}

}  // namespace xt
}  // namespace gcnb

// ---- host side ---------------------------------------------------------------------------------------
namespace gcnb {

// D[M x N] = X^T · W with X stored [K x M] (leading dimension ldx) and W stored [K x N] (ldw).
// Returns GCNB_OK when the product ran here, 1 when the shape is not this kernel's (caller falls back).
// colsum (optional): receives 1^T·W, the column sums of W, from the same pass (ones column in the X operand);
// returns 1 when that cannot be done here (no spare row in the 128-row tile).
int xw_tallk_stream_run(const float* X, int64_t ldx, const float* W, int64_t ldw, float* D, int64_t ldd, int64_t M,
                        int N, int64_t K, const float* bias, int epi, int precision, int n_sm, float* colsum,
                        cudaStream_t stream) {
  static const int enabled = getenv("GCNB_XW_TALLK_STREAM") ? atoi(getenv("GCNB_XW_TALLK_STREAM")) : 1;
  if (!enabled || precision > 1 || K < 4096 || K >= ((int64_t)1 << 31) || M < 1 || N < 1) return 1;
  // whole rows travel as 16-byte-granular bulk copies
  if (((uintptr_t)X & 15) || ((uintptr_t)W & 15) || (ldx % 4) != 0 || (ldw % 4) != 0 || (M % 4) != 0 || (N % 4) != 0) return 1;
  const bool direct = M <= 128 && N <= 256;
  const bool swapped = !direct && N <= 128 && M <= 256;
  if (!direct && !swapped) return 1;
  if (colsum && !(direct && M < 128)) return 1;
  xt::Params p;
  p.A = direct ? X : W; p.lda = direct ? ldx : ldw;
  p.B = direct ? W : X; p.ldb = direct ? ldw : ldx;
  const int Mp = direct ? (int)M : N, Np = direct ? N : (int)M;
  p.K = K; p.M = Mp; p.N = Np; p.mb = (Mp + 31) / 32; p.nb = (Np + 31) / 32;
  p.ones_col = colsum ? Mp : -1;
  p.M1 = colsum ? Mp + 1 : Mp;
  p.raw_stage = ((xt::KV * (Mp + Np) * 4) + 127) & ~127;
  p.raws = std::min(xt::RAW_MAX, (xt::SMEM_MAX - xt::CS * xt::CONV_STAGE - 2048) / p.raw_stage);
  if (p.raws < 2) return 1;
  const int smem = xt::CS * xt::CONV_STAGE + p.raws * p.raw_stage + 2048;
  const int64_t stages_total = (K + xt::KV - 1) / xt::KV;
  const int grid = (int)std::min<int64_t>(n_sm, stages_total);
  const int64_t st_per_cta = (stages_total + grid - 1) / grid;
  p.v_per_cta = st_per_cta * xt::KV;
  const char* chain_env = getenv("GCNB_XW_MAX_CHAIN");
  p.chain = std::max(1, chain_env ? atoi(chain_env) : 32);
  float* part = nullptr;
  GCNB_CHECK_CUDA(cudaMallocAsync(&part, (size_t)grid * p.M1 * Np * sizeof(float), stream));
  p.part = part;
  static std::atomic<uint64_t> configured{0};
  int dev = 0;
  GCNB_CHECK_CUDA(cudaGetDevice(&dev));
  const uint64_t bit = 1ull << (dev & 63);
  if (dev >= 64 || !(configured.load(std::memory_order_relaxed) & bit)) {
    GCNB_CHECK_CUDA(cudaFuncSetAttribute(xt::xt_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, xt::SMEM_MAX));
    GCNB_CHECK_CUDA(cudaFuncSetAttribute(xt::xt_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, xt::SMEM_MAX));
    configured.fetch_or(bit, std::memory_order_relaxed);
  }
  if (precision == 0) xt::xt_kernel<0><<<grid, xt::THREADS, smem, stream>>>(p);
  else xt::xt_kernel<1><<<grid, xt::THREADS, smem, stream>>>(p);
  int rc = GCNB_OK;
  if (cudaGetLastError() != cudaSuccess) rc = set_error(GCNB_ERR_CUDA, "xt_kernel launch failed");
  if (rc == GCNB_OK) {
    // CTAs past the end of K wrote nothing
    const int n_valid = (int)std::min<int64_t>(grid, (K + p.v_per_cta - 1) / p.v_per_cta);
    xt::xt_reduce_kernel<<<(unsigned)((p.M1 * Np / 4 + 127) / 128), 128, 0, stream>>>(
        part, n_valid, p.M1, Mp, Np, direct ? 0 : 1, D, ldd, (int)M, N, bias, epi, p.ones_col, colsum);
    if (cudaGetLastError() != cudaSuccess) rc = set_error(GCNB_ERR_CUDA, "xt_reduce_kernel launch failed");
  }
  cudaFreeAsync(part, stream);
  return rc;
}

}  // namespace gcnb
