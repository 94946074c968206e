// Dense feature transform for the layer shapes of the GCN epoch — D[M x N] = X[M x K] · op(W) (+bias)(ReLU)
// (dropout)(mask) with M = vertices (millions), K and N = feature widths (<= a few hundred) — as a
// persistent, warp-specialised tcgen05 kernel.  Replaces torch.mm(input, self.weight) of the reference
// layer (pygcn/gcn6.py:131, :191; gcn5.py:88, :136), the F.dropout that follows it (gcn6.py:246) and, in
// the backward pass, the dropout / ReLU masking of dH = dS·W^T (autograd of gcn6.py:245-246).
//
// Why a second kernel next to xw_tcgen05.cu: that one (one 4-warp CTA per 128 x BN tile, k-blocks of 16,
// every fp32 value converted to tf32 hi/lo by the same threads that wait for it, a CTA-wide barrier per
// k-block, per-thread row stores) ran at ~27 % of the HBM time of these shapes.
// Here one CTA per SM loops over its 128-row tiles, 20 warps with fixed roles and separate rings per stream:
//   warp 0      X producer: one TMA tensor-map load per k-chunk (128 rows x 32 floats, SWIZZLE_128B, rows past M
//               and columns past K zero-filled by the TMA unit) into a ring that takes all shared memory left
//               (3-9 chunks).  One producer thread streams at the HBM rate this way once >= 64 KB are in flight
//               (tools/tma_stream_probe.cu); 16-byte cp.async copies by four loader warps were slower
//   warp 2      W producer: TMA bulk copies of the prepared W chunks (hi | lo images, L2-resident), once per CTA
//               when they fit (128 KB for a 128-column tile, 64 KB for a 64-column one), else through a ring of 2
//   warps 4-11  converters: split every raw chunk into hi (the fp32 word with its 13 low mantissa bits cleared) and
//               lo = tf32(x - hi) and write both into TENSOR memory (tcgen05.st, thread = row): the MMA takes its
//               A operand from there, only W comes from shared memory
//   warp 1      MMA issuer: the whole warp walks the loop, one elected lane issues tcgen05.mma kind::tf32 (A from
//               tensor memory, B = W from shared memory), accumulator in TMEM, two accumulator buffers so the
//               epilogue of tile i overlaps the MMAs of tile i+1; tcgen05.commit frees the rings / publishes the
//               accumulator
//   warps 12-19 epilogue: tcgen05.ld, bias / ReLU / dropout / mask with thread = row (one warp-uniform branch per
//               piece), one 32-bit word of mask bits per 32 outputs, then a transpose through a private
//               shared-memory tile so that every store instruction writes full 128-byte lines
//   * x·w = hi_x·hi_w + hi_x·lo_w + lo_x·hi_w (three MMAs per k-step of 8), error ~1e-6 of fp32 like the
//     first kernel (W is split with round-to-nearest by the prep kernel; |lo_x·lo_w| <= 2^-21 |x·w|).
//   * fused epilogues (GCNB_EPI_*): bias, ReLU, DROPOUT (keep bits from a counter hash of (seed, group of 4
//     elements), p quantised to 1/256; out = keep ? out / (1-p) : 0), optional output of the bit mask
//     [out > 0] (1 bit per element), MASK (out = bit ? out / (1-p) : 0 — the backward of ReLU followed by
//     dropout costs 1/32 of a word per element instead of re-reading the forward activation).
// Used by gcnb_xw / gcnb_xw_ex when X is not transposed, 16-byte aligned with ldx % 4 == 0, M >= 1024 and
// K <= 4096, precision 0 or 1; everything else stays on xw_tcgen05.cu.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>
#include <atomic>

#include "common.h"

namespace gcnb {
namespace xs {

constexpr int BM = 128;          // rows per tile == UMMA M
constexpr int KC = 32;           // floats per k-chunk: one 128-byte swizzle row
constexpr int CHUNK_A = BM * 128;                 // bytes of one A chunk (raw or lo)
constexpr int STG_PITCH = 144;   // bytes per row of the epilogue's transpose tile (32 floats + 16 pad)
constexpr int STG_BYTES = 32 * STG_PITCH;
constexpr int SMEM_MAX = 232448; // 227 KB of dynamic shared memory per CTA

constexpr int NE = 8;            // epilogue warps (two per TMEM lane quadrant)
constexpr int NCV = 8;           // converter warps
constexpr int THREADS = (4 + NCV + NE) * 32;
constexpr int RA_MAX = 12;       // most X chunks in flight (barrier slots)
__host__ __device__ constexpr int w_chunk(int bn) { return 2 * bn * 128; }      // hi | lo
// shared memory: X ring (ra chunks) | W (w_slots chunks: 2 when streamed, nkc when resident) |
// epilogue transpose tiles | barriers; + 1 KB alignment slack
// tensor memory (512 columns): two accumulators of BN columns | ring of NA converted X chunks, 64 columns each
// (hi: 32 k-values | lo: 32)
constexpr int NA_MAX = 4;        // most converted chunks in flight
constexpr int A_COLS = 2 * KC;   // TMEM columns of one converted chunk
constexpr int kTraceLen = 512;
__host__ __device__ constexpr int fixed_bytes(int bn, int w_slots) {
  return w_slots * w_chunk(bn) + NE * STG_BYTES + 1024 /* barriers */ + 1024 /* alignment */;
}
__host__ __device__ constexpr int ring_a(int bn, int w_slots) {
  return (SMEM_MAX - fixed_bytes(bn, w_slots)) / CHUNK_A < RA_MAX ? (SMEM_MAX - fixed_bytes(bn, w_slots)) / CHUNK_A : RA_MAX;
}

struct Params {
  const uint8_t* w_img;      // [ntiles][nkc] { hi[bn x 128 B] | lo[bn x 128 B] }, swizzled like the A chunk
  float* D; int64_t ldd;
  const uint32_t* mask_in;   // GCNB_EPI_MASK: bit (row, col) = word[row * ldm + col / 32] >> (col % 32)
  uint32_t* mask_out;        // optional: bit = (out > 0) after the epilogue
  int64_t ldm;               // words per row of the bit masks
  const float* bias;
  int64_t M; int N; int K;
  int nkc, ntiles;
  int ra;                    // X chunks in flight (ring slots)
  int na;                    // converted chunks in flight in tensor memory (2..NA_MAX)
  int groups;                // resident-W mode: CTAs per n-tile (CTA b works on n-tile b % ntiles, m-tiles b / ntiles + i * groups)
  int64_t mtiles;
  int epi;
  float scale;               // 1/(1-p) for DROPOUT and MASK (p quantised to 1/256)
  uint32_t keep_thresh;      // DROPOUT: keep when the element's hash byte >= thresh (0..255)
  uint32_t seed;
  int d_vec;                 // 16-byte stores of D allowed
  unsigned long long* trace; // GCNB_XS_TRACE: clock64() stamps of CTA 0, [8 roles][kTraceLen], diagnosis only
  int dbg;                   // GCNB_XS_DEBUG bits (diagnosis only, results wrong): 1 MMA hi·hi only, 2 converters skip the
                             // low part, 4 epilogue skips the global stores, 8 every X load reads tile 0 (L2 hits)
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
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// shared-memory matrix descriptor: K-major, SWIZZLE_128B (layout type 2), 8-row groups 1024 bytes apart
__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3fff);
  d |= (uint64_t)1 << 16;                       // leading byte offset: unused for swizzled K-major
  d |= (uint64_t)(1024 >> 4) << 32;             // stride byte offset
  d |= (uint64_t)1 << 46;                       // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                       // SWIZZLE_128B
  return d;
}
// instruction descriptor: D fp32, A/B TF32, both K-major, M = 128, N = n
__host__ __device__ constexpr uint32_t make_idesc(int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
}
__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
// A operand in tensor memory (lane = row, one 32-bit column per k), B from shared memory
__device__ __forceinline__ void mma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
        "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
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
// counter hash for the dropout keep bits: one 32-bit word per group of 4 consecutive columns of a row
// (group index = row * ceil(N/4) + col/4), one byte per element
__host__ __device__ __forceinline__ uint32_t mix32(uint64_t idx, uint32_t seed) {
  uint32_t x = (uint32_t)idx ^ seed, y = (uint32_t)(idx >> 32) + 0x9e3779b9u * (seed | 1u);
  x *= 0x85ebca6bu; x ^= x >> 13; x += y; x *= 0xc2b2ae35u; x ^= x >> 16; x *= 0x27d4eb2fu; x ^= x >> 15;
  return x;
}
__host__ __device__ __forceinline__ uint32_t drop_thresh(float p) {   // p quantised to 1/256, < 1
  int t = (int)(p * 256.f + 0.5f);
  return (uint32_t)(t < 0 ? 0 : (t > 255 ? 255 : t));
}

}  // namespace xs
}  // namespace gcnb

namespace gcnb {
namespace xs {

// WRES: the W chunks of the CTA's n-tile stay resident in shared memory for the whole kernel (loaded once);
// otherwise they stream through a ring of 2 per k-chunk (large K).  With resident W a CTA keeps ONE n-tile
// and the CTAs b, b+1, .. b+ntiles-1 walk the same m-tiles at the same time, so that an X chunk comes out of
// HBM once and out of L2 for the others.
template <int BN, int MODE, bool WRES>   // MODE 0: 3xTF32 (fp32 parity), 1: single-pass TF32
__global__ void __launch_bounds__(THREADS, 1)
xs_kernel(const __grid_constant__ CUtensorMap tmap_x, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  const int RA = p.ra;                            // X chunks in flight
  constexpr int NEG = NE / 4;                     // epilogue warps per TMEM lane quadrant
  constexpr int W_CHUNK = w_chunk(BN);
  constexpr uint32_t IDESC = make_idesc(BN);
  constexpr int TMEM_COLS = 512;                  // two accumulator buffers (2 x BN columns) + the converted X chunks
  const int w_slots = WRES ? p.nkc : 2;
  // 1024-byte alignment (of the shared-space address) for the 128-byte swizzle atoms
  uint8_t* smem = smem_raw + ((1024u - (smem_addr(smem_raw) & 1023u)) & 1023u);
  uint8_t* ring = smem;                            // RA x 16 KB raw X chunks
  const int NA = p.na;
  uint8_t* wbuf = ring + RA * CHUNK_A;             // W chunks
  uint8_t* stg = wbuf + w_slots * W_CHUNK;         // NE transpose tiles
  uint64_t* bars = reinterpret_cast<uint64_t*>(stg + NE * STG_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * RA_MAX + 8 + 2 * NA_MAX);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t bar0 = smem_addr(bars);
  auto FULLA = [&](int s) { return bar0 + 8u * (uint32_t)s; };
  auto EMPTYA = [&](int s) { return bar0 + 8u * (uint32_t)(RA_MAX + s); };
  auto FULLW = [&](int s) { return bar0 + 8u * (uint32_t)(2 * RA_MAX + s); };
  auto EMPTYW = [&](int s) { return bar0 + 8u * (uint32_t)(2 * RA_MAX + 2 + s); };
  auto TFULL = [&](int b) { return bar0 + 8u * (uint32_t)(2 * RA_MAX + 4 + b); };
  auto TEMPTY = [&](int b) { return bar0 + 8u * (uint32_t)(2 * RA_MAX + 6 + b); };
  auto CONV = [&](int s) { return bar0 + 8u * (uint32_t)(2 * RA_MAX + 8 + s); };
  auto AEMPTY = [&](int s) { return bar0 + 8u * (uint32_t)(2 * RA_MAX + 8 + NA_MAX + s); };
  // work of this CTA: units u = u0, u0 + ustep, ... of (m-tile, n-tile) pairs
  //   streamed W: unit index t over all pairs, n-tile fastest;   resident W: fixed n-tile, m-tiles of its group
  const int64_t n_tiles = p.mtiles * p.ntiles;
  const int my_nt = WRES ? (int)(blockIdx.x % p.ntiles) : 0;
  const int64_t u0 = WRES ? (int64_t)(blockIdx.x / p.ntiles) : (int64_t)blockIdx.x;
  const int64_t ustep = WRES ? (int64_t)p.groups : (int64_t)gridDim.x;
  const int64_t uend = WRES ? p.mtiles : n_tiles;
  auto unit_nt = [&](int64_t u) { return WRES ? my_nt : (int)(u % p.ntiles); };
  auto unit_m0 = [&](int64_t u) { return (WRES ? u : u / p.ntiles) * BM; };

  if (tid == 0) {
    for (int s = 0; s < RA_MAX; ++s) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(FULLA(s)));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(EMPTYA(s)), "r"(NCV * 32));
    }
    for (int b = 0; b < NA_MAX; ++b) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(CONV(b)), "r"(NCV * 32));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(AEMPTY(b)));
    }
    for (int b = 0; b < 2; ++b) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(FULLW(b)));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(EMPTYW(b)));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(TFULL(b)));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(TEMPTY(b)), "r"(NE * 32));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_addr(tmem_slot)), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  auto stamp = [&](int role, uint32_t idx) {
    if (p.trace != nullptr && blockIdx.x == 0 && idx < (uint32_t)kTraceLen) p.trace[role * kTraceLen + idx] = (unsigned long long)clock64();
  };

  if (warp == 0) {
    // ---------------- X producer: one TMA tensor-map load per k-chunk ----------------
    // (16-byte cp.async copies by four loader warps into the same layout were tried instead — no tensor map, any
    // leading dimension — and were slower: layer-1 product 2.72 vs 2.43 ms, profiles/r02_k_xw_shapes.jsonl.)
    if (lane == 0) {
      uint32_t it = 0;
      int s = 0;                // ring position and phase kept as counters: an integer division per chunk is a few
      uint32_t ph = 0;          // dozen dependent instructions of a single warp, i.e. hundreds of cycles
      for (int64_t t = u0; t < uend; t += ustep) {
        const int64_t m0 = unit_m0(t);
        for (int kc = 0; kc < p.nkc; ++kc, ++it) {
          mbar_wait(EMPTYA(s), ph ^ 1);
          const uint32_t full = FULLA(s);
          asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(full), "r"((uint32_t)CHUNK_A) : "memory");
          asm volatile(
              "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
              ::"r"(smem_addr(ring + s * CHUNK_A)), "l"(&tmap_x), "r"(kc * KC), "r"((p.dbg & 8) ? 0 : (int)m0), "r"(full) : "memory");
          stamp(0, it);
          if (++s == RA) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 2) {
    // ---------------- W producer ----------------
    if (lane == 0) {
      if constexpr (WRES) {
        if (u0 < uend) {   // every chunk of this CTA's n-tile, once
          const uint32_t full = FULLW(0);
          asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(full), "r"((uint32_t)(p.nkc * W_CHUNK)) : "memory");
          for (int kc = 0; kc < p.nkc; ++kc) {
            const uint8_t* src_w = p.w_img + ((int64_t)my_nt * p.nkc + kc) * W_CHUNK;
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(smem_addr(wbuf + kc * W_CHUNK)), "l"(src_w), "r"((uint32_t)W_CHUNK), "r"(full) : "memory");
          }
        }
      } else {
        uint32_t it = 0;
        for (int64_t t = u0; t < uend; t += ustep) {
          const int nt = unit_nt(t);
          for (int kc = 0; kc < p.nkc; ++kc, ++it) {
            const int s = it & 1;
            mbar_wait(EMPTYW(s), ((it >> 1) & 1) ^ 1);
            const uint32_t full = FULLW(s);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(full), "r"((uint32_t)W_CHUNK) : "memory");
            const uint8_t* src_w = p.w_img + ((int64_t)nt * p.nkc + kc) * W_CHUNK;
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(smem_addr(wbuf + s * W_CHUNK)), "l"(src_w), "r"((uint32_t)W_CHUNK), "r"(full) : "memory");
          }
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer ----------------
    // The whole warp walks the loop (uniform control flow, operands in uniform registers) and one elected lane
    // issues.  Under `if (lane == 0)` the compiler wraps every tcgen05.mma in a uniformising loop (ELECT, five
    // R2UR.BROADCAST, BRA.U.ANY): ~85 cycles per MMA, 1000 per chunk — as long as the chunk's load or conversion
    // (clock64 timeline of the roles, profiles/r02_t_xs_trace_report.txt).
    {
      // This warp's own instruction stream is what the tensor pipe waits for (a single warp retires a dependent
      // instruction every ~4+ cycles): ring positions are counters, the W descriptors are one add away from a base
      // descriptor (16-byte units in the low word), nothing is recomputed per MMA (profiles/r02_w_xs_trace_report.txt:
      // ~1000 cycles per chunk of 10 MMAs before this).
      uint32_t it = 0, tl = 0;
      int sn = 0;
      uint32_t ph_n = 0;
      const uint64_t dw0 = make_desc_sw128(smem_addr(wbuf));
      if constexpr (WRES) {
        if (u0 < uend) mbar_wait(FULLW(0), 0);
      }
      for (int64_t t = u0; t < uend; t += ustep, ++tl) {
        const uint32_t buf = tl & 1;
        mbar_wait(TEMPTY(buf), ((tl >> 1) & 1) ^ 1);       // the epilogue has drained this accumulator
        const uint32_t tmem_d = tmem_base + buf * BN;
        for (int kc = 0; kc < p.nkc; ++kc, ++it) {
          const int sw = it & 1;
          if (lane == 0) stamp(1, it);
          if constexpr (!WRES) mbar_wait(FULLW(sw), (it >> 1) & 1);
          mbar_wait(CONV(sn), ph_n);                   // the converters have put hi | lo of this chunk into tensor memory
          if (lane == 0) stamp(2, it);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          if (lane == 0) stamp(4, it);
          const uint32_t a_hi = tmem_base + 2 * BN + sn * A_COLS, a_lo = a_hi + KC;
          const uint64_t d_hi = dw0 + (uint64_t)((WRES ? kc : sw) * (W_CHUNK >> 4)), d_lo = d_hi + (uint64_t)((BN * 128) >> 4);
          // k-steps of 8 that hold columns < K (the tail chunk of K = 100 has one, not four; the TMA unit zero-fills
          // the columns past K inside it)
          const int ksteps = min(KC / 8, (p.K - kc * KC + 7) >> 3);
          __syncwarp();
          if (elect_one()) {
#pragma unroll
            for (int ks = 0; ks < KC / 8; ++ks) {     // K = 8 tf32 per MMA: 8 TMEM columns of A, 32 bytes of W's swizzle row
              if (ks < ksteps) {
                const uint32_t acc = (kc > 0 || ks > 0) ? 1u : 0u;
                mma_tf32_ts(tmem_d, a_hi + ks * 8, d_hi + (uint64_t)(ks * 2), IDESC, acc);
                if (MODE == 0 && !(p.dbg & 1)) {
                  mma_tf32_ts(tmem_d, a_hi + ks * 8, d_lo + (uint64_t)(ks * 2), IDESC, 1u);
                  mma_tf32_ts(tmem_d, a_lo + ks * 8, d_hi + (uint64_t)(ks * 2), IDESC, 1u);
                }
              }
            }
            if constexpr (!WRES) commit(EMPTYW(sw));
            commit(AEMPTY(sn));                        // the chunk's TMEM columns are free once these MMAs have read them
            if (kc == p.nkc - 1) commit(TFULL(buf));   // accumulator complete
          }
          __syncwarp();
          if (lane == 0) stamp(5, it);
          if (++sn == NA) { sn = 0; ph_n ^= 1; }
        }
      }
    }
  } else if (warp >= 4 && warp < 4 + NCV) {
    // ---------------- converters: raw fp32 chunk in shared memory -> hi | lo in TENSOR memory ----------------
    // hi = the fp32 word with its 13 low mantissa bits cleared, lo = tf32(x - hi); thread = row (TMEM lane), two
    // warps per lane quadrant, 16 of the chunk's 32 k-values each.  With the A operand in tensor memory an MMA reads
    // only W's 2-4 KB from shared memory instead of 6-8 KB (the tensor pipe was pacing itself on those reads:
    // profiles/r02_u_xs_trace_report.txt), the in-place hi store and the low-part ring (32 KB) are gone, and the
    // X ring is two chunks deeper.  Reading one's own 128-byte row is conflict-free because of the swizzle
    // (the 8 lanes of a quarter-warp hit 8 different 16-byte columns).
    {
      const int q = warp & 3;                        // TMEM lanes 32q .. 32q+31 belong to this warp
      const int half = (warp - 4) >> 2;              // k-values 16*half .. 16*half+15 of the chunk
      const int row = q * 32 + lane;
      uint32_t it = 0;
      int sa = 0, sn = 0;
      uint32_t ph_a = 0, ph_n = 0;
      for (int64_t t = u0; t < uend; t += ustep) {
        for (int kc = 0; kc < p.nkc; ++kc, ++it) {
          mbar_wait(FULLA(sa), ph_a);
          mbar_wait(AEMPTY(sn), ph_n ^ 1);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          if (!(p.dbg & 2)) {
            const uint8_t* src = ring + sa * CHUNK_A + row * 128;
            uint32_t hi[16], lo[16];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              const uint4 v = *reinterpret_cast<const uint4*>(src + (((half * 4 + c) ^ (row & 7)) << 4));
              const uint32_t x[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                hi[4 * c + e] = x[e] & 0xffffe000u;
                if constexpr (MODE == 0) lo[4 * c + e] = to_tf32(__uint_as_float(x[e]) - __uint_as_float(x[e] & 0xffffe000u));
              }
            }
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + 2 * BN + sn * A_COLS + half * 16;
            tmem_st16(taddr, hi);
            if constexpr (MODE == 0) tmem_st16(taddr + KC, lo);
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
          }
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          mbar_arrive(CONV(sn));
          mbar_arrive(EMPTYA(sa));                     // this thread's reads of the raw chunk are done
          if (warp == 3 + NCV && lane == 31) stamp(3, it);
          if (++sa == RA) { sa = 0; ph_a ^= 1; }
          if (++sn == NA) { sn = 0; ph_n ^= 1; }
        }
      }
    }
  } else if (warp >= 4 + NCV) {
    // ---------------- epilogue ----------------
    const int q = warp & 3;                          // TMEM lanes 32q .. 32q+31 belong to this warp
    const int eg = (warp - 4 - NCV) >> 2;            // which of the NEG warps of this quadrant
    uint8_t* my = stg + (warp - 4 - NCV) * STG_BYTES;
    const int64_t n4 = (p.N + 3) / 4;                // dropout hash groups per row
    uint32_t tl = 0;
    for (int64_t t = u0; t < uend; t += ustep, ++tl) {
      const uint32_t buf = tl & 1;
      const int nt = unit_nt(t);
      const int64_t m0 = unit_m0(t) + q * 32;
      const int n0 = nt * BN;
      const int64_t myrow = m0 + lane;
      const uint32_t taddr0 = tmem_base + buf * BN + ((uint32_t)(q * 32) << 16);
      bool released = false;
#pragma unroll 1
      for (int c0 = eg * 32; c0 < BN; c0 += NEG * 32) {
        if (n0 + c0 >= p.N) break;                   // warp-uniform
        const int colb = n0 + c0;
        const int epi = p.epi;
        // operands of the elementwise part, requested before the accumulator is waited for (their latency — DRAM
        // for the mask word — hides behind the tcgen05.ld): lane j holds the bias of column colb + j
        float bias_j = 0.f;
        if ((epi & GCNB_EPI_BIAS) && colb + lane < p.N) bias_j = __ldg(p.bias + colb + lane);
        uint32_t min_word = 0xffffffffu;
        if ((epi & GCNB_EPI_MASK) && myrow < p.M) min_word = p.mask_in[myrow * p.ldm + (colb >> 5)];
        if (c0 == eg * 32) {
          mbar_wait(TFULL(buf), (tl >> 1) & 1);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          if (warp == 4 + NCV && lane == 0) stamp(6, tl);
        }
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
        if (c0 + NEG * 32 >= BN || n0 + c0 + NEG * 32 >= p.N) {
          // this warp's last read of the accumulator: hand the buffer back before the elementwise work
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          mbar_arrive(TEMPTY(buf));
          released = true;
        }
        // ---- elementwise epilogue, thread = row, 32 consecutive columns in registers.  One warp-uniform branch
        // per PIECE, each piece a straight run over the 32 values: with the flags tested per element the loop was
        // ~1200 predicated instructions per slice and the epilogue warps, not memory, set the pace of the kernel
        // (profiles/r02_q_xs_debug.jsonl: 1.39 ms with loads from L2, one MMA pass and no stores) ----
        if (epi & GCNB_EPI_BIAS) {
          __syncwarp();
          reinterpret_cast<float*>(my)[lane] = bias_j;            // 128 bytes of the transpose tile, before it is used
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 b4 = *reinterpret_cast<const float4*>(my + j * 4);   // broadcast read
            r[j] = __float_as_uint(__uint_as_float(r[j]) + b4.x);
            r[j + 1] = __float_as_uint(__uint_as_float(r[j + 1]) + b4.y);
            r[j + 2] = __float_as_uint(__uint_as_float(r[j + 2]) + b4.z);
            r[j + 3] = __float_as_uint(__uint_as_float(r[j + 3]) + b4.w);
          }
        }
        if (epi & GCNB_EPI_RELU) {
#pragma unroll
          for (int j = 0; j < 32; ++j) r[j] = __float_as_uint(fmaxf(__uint_as_float(r[j]), 0.f));
        }
        if (epi & GCNB_EPI_DROPOUT) {
          const uint64_t g0 = (uint64_t)myrow * (uint64_t)n4 + (uint64_t)(colb >> 2);
#pragma unroll
          for (int g4 = 0; g4 < 8; ++g4) {
            const uint32_t h = mix32(g0 + (uint64_t)g4, p.seed);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int j = g4 * 4 + e;
              r[j] = ((h >> (8 * e)) & 0xffu) >= p.keep_thresh ? __float_as_uint(__uint_as_float(r[j]) * p.scale) : 0u;
            }
          }
        }
        if (epi & GCNB_EPI_MASK) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            r[j] = ((min_word >> j) & 1u) ? __float_as_uint(__uint_as_float(r[j]) * p.scale) : 0u;
        }
        if (p.mask_out != nullptr) {
          uint32_t out_word = 0;
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (__uint_as_float(r[j]) > 0.f) out_word |= 1u << j;
          if (colb + 32 > p.N) out_word &= (1u << (p.N - colb)) - 1u;   // columns past N carry no bit
          if (myrow < p.M) p.mask_out[myrow * p.ldm + (colb >> 5)] = out_word;
        }
        // ---- transpose through the private tile: 16-byte chunks, pitch 144 (conflict-free) ----
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(my + lane * STG_PITCH + j * 16) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
        __syncwarp();
        // read back 4 rows x 128 bytes per instruction: full output lines
        const int ch = lane & 7;
        const int col = colb + ch * 4;
        if (p.dbg & 4) continue;
        if (p.d_vec && colb + 32 <= p.N && m0 + 32 <= p.M) {        // whole slice inside the matrix (warp-uniform)
          float* dp = p.D + (m0 + (lane >> 3)) * p.ldd + col;
#pragma unroll
          for (int i = 0; i < 8; ++i)
            *reinterpret_cast<float4*>(dp + (int64_t)(i * 4) * p.ldd) =
                *reinterpret_cast<const float4*>(my + (i * 4 + (lane >> 3)) * STG_PITCH + ch * 16);
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int rr = i * 4 + (lane >> 3);
            const int64_t row = m0 + rr;
            const float4 v = *reinterpret_cast<const float4*>(my + rr * STG_PITCH + ch * 16);
            if (row < p.M && col < p.N) {
              float* dp = p.D + row * p.ldd + col;
              if (p.d_vec && col + 4 <= p.N) {
                *reinterpret_cast<float4*>(dp) = v;
              } else {
                const float o[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int e = 0; e < 4; ++e)
                  if (col + e < p.N) dp[e] = o[e];
              }
            }
          }
        }
      }
      if (warp == 4 + NCV && lane == 0) stamp(7, tl);
      if (!released) {   // this warp had no column group in the tile (N small): still a party of the barrier
        mbar_wait(TFULL(buf), (tl >> 1) & 1);
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        mbar_arrive(TEMPTY(buf));
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
}

// W -> per (n-tile, k-chunk) images { hi | lo }, each bn rows (n) x 128 bytes (32 k), 16-byte chunks XOR-swizzled
// with the row like the TMA's SWIZZLE_128B; round-to-nearest split.  One thread per (n, k).
__global__ void __launch_bounds__(256)
xs_prep_w_kernel(const float* __restrict__ W, int64_t ldw, int trans_w, int N, int K, int bn, int nkc, int ntiles,
                 uint8_t* __restrict__ img) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t kpad = (int64_t)nkc * KC, npad = (int64_t)ntiles * bn;
  if (idx >= kpad * npad) return;
  int k, n;   // fastest index follows the contiguous direction of the stored matrix
  if (trans_w) { k = (int)(idx % kpad); n = (int)(idx / kpad); }
  else         { n = (int)(idx % npad); k = (int)(idx / npad); }
  float v = 0.f;
  if (n < N && k < K) v = trans_w ? W[(int64_t)n * ldw + k] : W[(int64_t)k * ldw + n];
  const int nt = n / bn, row = n % bn, kc = k / KC, kk = k % KC;
  uint8_t* tile = img + ((int64_t)nt * nkc + kc) * (2 * (int64_t)bn * 128);
  const int off = row * 128 + ((((kk >> 2) ^ (row & 7)) & 7) << 4) + (kk & 3) * 4;
  const uint32_t hi = to_tf32(v);
  *reinterpret_cast<uint32_t*>(tile + off) = hi;
  *reinterpret_cast<uint32_t*>(tile + (int64_t)bn * 128 + off) = to_tf32(v - __uint_as_float(hi));
}

}  // namespace xs
}  // namespace gcnb

// ---- host side ---------------------------------------------------------------------------------------
namespace gcnb {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled() {
  static std::atomic<EncodeTiledFn> fn{nullptr};
  EncodeTiledFn f = fn.load(std::memory_order_acquire);
  if (!f) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      f = (EncodeTiledFn)ptr;
    fn.store(f, std::memory_order_release);
  }
  return f;
}

template <int BN, int MODE, bool WRES>
static int xs_launch(const CUtensorMap& tmap, const xs::Params& p, int grid, int smem, cudaStream_t stream) {
  static std::atomic<uint64_t> configured{0};   // per-device opt-in to > 48 KB of dynamic shared memory
  int dev = 0;
  GCNB_CHECK_CUDA(cudaGetDevice(&dev));
  const uint64_t bit = 1ull << (dev & 63);
  if (dev >= 64 || !(configured.load(std::memory_order_relaxed) & bit)) {
    GCNB_CHECK_CUDA(cudaFuncSetAttribute(xs::xs_kernel<BN, MODE, WRES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         xs::SMEM_MAX));
    configured.fetch_or(bit, std::memory_order_relaxed);
  }
  xs::xs_kernel<BN, MODE, WRES><<<grid, xs::THREADS, smem, stream>>>(tmap, p);
  GCNB_CHECK_CUDA(cudaGetLastError());
  return GCNB_OK;
}

template <int MODE>
static int xs_dispatch(int bn, bool wres, const CUtensorMap& tmap, const xs::Params& p, int grid, int smem, cudaStream_t stream) {
  if (bn == 64) return wres ? xs_launch<64, MODE, true>(tmap, p, grid, smem, stream) : xs_launch<64, MODE, false>(tmap, p, grid, smem, stream);
  return wres ? xs_launch<128, MODE, true>(tmap, p, grid, smem, stream) : xs_launch<128, MODE, false>(tmap, p, grid, smem, stream);
}

// Returns GCNB_OK when the product ran here, 1 when the shape is not this kernel's (caller falls back).
int xw_stream_run(const float* X, int64_t ldx, const float* W, int64_t ldw, int trans_w, float* D, int64_t ldd,
                  int64_t M, int N, int K, const float* bias, int epi, const uint32_t* mask_in, uint32_t* mask_out,
                  int64_t ldm, float drop_p, uint32_t seed, int precision, int n_sm, cudaStream_t stream) {
  static const int enabled = getenv("GCNB_XW_STREAM") ? atoi(getenv("GCNB_XW_STREAM")) : 1;
  // GCNB_XW_WRES: 1 (default) keeps the W chunks of a CTA's n-tile resident when they fit 64 KB, 0 always streams them
  static const int wres_ok = getenv("GCNB_XW_WRES") ? atoi(getenv("GCNB_XW_WRES")) : 1;
  if (!enabled || precision > 1 || M < 1024 || K < 1 || K > 4096 || N < 1) return 1;
  if (((uintptr_t)X & 15) || (ldx % 4) != 0 || M >= ((int64_t)1 << 31)) return 1;
  const int nkc = (K + xs::KC - 1) / xs::KC;
  // tile width and W placement: resident W (<= 64 KB: all k-chunks of one n-tile) with the widest tile that
  // allows it, else streamed W chunks
  int bn = N <= 64 ? 64 : 128;
  bool wres = false;
  // GCNB_XS_WRES_KB: most shared memory (KB) the resident W image of a 128-column tile may take; beyond it the tile
  // is 64 columns wide (half the image, the X chunks read twice as often)
  static const int wres_kb = getenv("GCNB_XS_WRES_KB") ? atoi(getenv("GCNB_XS_WRES_KB")) : 128;
  if (wres_ok) {
    if (N > 64 && nkc * xs::w_chunk(128) <= wres_kb * 1024) { bn = 128; wres = true; }
    else if (nkc * xs::w_chunk(64) <= 65536) { bn = 64; wres = true; }
  }
  int ntiles = (N + bn - 1) / bn;
  if (wres && ntiles > n_sm) { wres = false; bn = N <= 64 ? 64 : 128; ntiles = (N + bn - 1) / bn; }
  const int w_slots = wres ? nkc : 2;
  // ring depths: GCNB_XS_NA converted chunks in tensor memory (2..4, default 4), all of shared memory that is left
  // is the X ring (GCNB_XS_RA caps it)
  static const int na_env = getenv("GCNB_XS_NA") ? atoi(getenv("GCNB_XS_NA")) : xs::NA_MAX;
  static const int ra_env = getenv("GCNB_XS_RA") ? atoi(getenv("GCNB_XS_RA")) : xs::RA_MAX;
  const int na_slots = std::max(2, std::min(xs::NA_MAX, na_env));
  const int ra = std::min(std::max(2, ra_env), xs::ring_a(bn, w_slots));
  if (ra < 2) return 1;
  const int smem = xs::fixed_bytes(bn, w_slots) + ra * xs::CHUNK_A;
  EncodeTiledFn enc = encode_tiled();
  if (!enc) return 1;
  CUtensorMap tmap;
  const cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)M};
  const cuuint64_t gstride[1] = {(cuuint64_t)ldx * 4};
  const cuuint32_t box[2] = {(cuuint32_t)xs::KC, (cuuint32_t)xs::BM};
  const cuuint32_t estr[2] = {1, 1};
  if (enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)X, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return 1;
  uint8_t* img = nullptr;
  const size_t img_bytes = (size_t)ntiles * nkc * xs::w_chunk(bn);
  GCNB_CHECK_CUDA(cudaMallocAsync(&img, img_bytes, stream));
  {
    const int64_t total = (int64_t)ntiles * bn * nkc * xs::KC;
    xs::xs_prep_w_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(W, ldw, trans_w, N, K, bn, nkc, ntiles, img);
    if (cudaGetLastError() != cudaSuccess) { cudaFreeAsync(img, stream); return set_error(GCNB_ERR_CUDA, "xs_prep_w_kernel launch failed"); }
  }
  xs::Params p;
  p.w_img = img; p.D = D; p.ldd = ldd; p.mask_in = mask_in; p.mask_out = mask_out; p.ldm = ldm; p.bias = bias;
  p.M = M; p.N = N; p.K = K; p.nkc = nkc; p.ntiles = ntiles; p.mtiles = (M + xs::BM - 1) / xs::BM;
  p.ra = ra; p.na = na_slots;
  p.epi = epi;
  p.keep_thresh = xs::drop_thresh(drop_p);
  p.scale = (epi & (GCNB_EPI_DROPOUT | GCNB_EPI_MASK)) ? 256.f / (256.f - (float)p.keep_thresh) : 1.f;
  p.seed = seed;
  p.d_vec = (((uintptr_t)D & 15) == 0) && (ldd % 4 == 0) && (N % 4 == 0);
  static const int dbg_env = getenv("GCNB_XS_DEBUG") ? atoi(getenv("GCNB_XS_DEBUG")) : 0;
  p.dbg = dbg_env;
  // GCNB_XS_TRACE=<file>: clock64() stamps of CTA 0's roles (producer issue, converter waits / arrive, MMA issue /
  // commit, epilogue start / end) for the first kTraceLen chunks / tiles, appended to <file> after a device
  // synchronisation.  Diagnosis only (tools/README.md).
  static const char* trace_path = getenv("GCNB_XS_TRACE");
  p.trace = nullptr;
  if (trace_path) {
    GCNB_CHECK_CUDA(cudaMalloc(&p.trace, 8 * xs::kTraceLen * sizeof(unsigned long long)));
    GCNB_CHECK_CUDA(cudaMemset(p.trace, 0, 8 * xs::kTraceLen * sizeof(unsigned long long)));
  }
  int grid;
  if (wres) {
    p.groups = (int)std::max<int64_t>(1, std::min<int64_t>(p.mtiles, n_sm / ntiles));
    grid = p.groups * ntiles;
  } else {
    p.groups = 0;
    grid = (int)std::min<int64_t>(p.mtiles * ntiles, n_sm);
  }
  const int rc = precision == 0 ? xs_dispatch<0>(bn, wres, tmap, p, grid, smem, stream)
                                : xs_dispatch<1>(bn, wres, tmap, p, grid, smem, stream);
  cudaFreeAsync(img, stream);
  if (p.trace) {
    static unsigned long long host[8 * xs::kTraceLen];
    if (cudaDeviceSynchronize() == cudaSuccess &&
        cudaMemcpy(host, p.trace, sizeof(host), cudaMemcpyDeviceToHost) == cudaSuccess) {
      if (FILE* f = fopen(trace_path, "a")) {
        fprintf(f, "# M=%lld N=%d K=%d bn=%d wres=%d nkc=%d ra=%d na=%d dbg=%d precision=%d\n", (long long)M, N, K, bn, (int)wres, nkc, ra,
                na_slots, p.dbg, precision);
        for (int r = 0; r < 8; ++r) {
          fprintf(f, "%d", r);
          for (int i = 0; i < xs::kTraceLen; ++i) fprintf(f, " %llu", host[r * xs::kTraceLen + i]);
          fprintf(f, "\n");
        }
        fclose(f);
      }
    }
    cudaFree(p.trace);
  }
  return rc;
}

// the keep decision of GCNB_EPI_DROPOUT, exported for tests: 1 when element (row, col) of an M x N output is kept
extern "C" int gcnb_dropout_keeps(int64_t row, int col, int N, float p, uint32_t seed) {
  const int64_t n4 = (N + 3) / 4;
  const uint32_t h = xs::mix32((uint64_t)row * (uint64_t)n4 + (uint64_t)(col >> 2), seed);
  return ((h >> (8 * (col & 3))) & 0xffu) >= xs::drop_thresh(p) ? 1 : 0;
}

}  // namespace gcnb
