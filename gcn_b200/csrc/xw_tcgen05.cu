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
#include <cuda_bf16.h>

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
__host__ __device__ constexpr int stage_bytes(int bn) { return 2 * A_PART + 2 * b_part_bytes(bn); }
__host__ __device__ constexpr int smem_bytes(int bn) {
  return 2 * stage_bytes(bn) + RAW_STAGES * RAW_BYTES + 64;
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
  const uint8_t* w_img;                       // prepared B operand image
  float* D; int64_t ldd;                      // output (or partials when splits > 1)
  int64_t M; int N; int K;
  int nkb;                                    // k-blocks in total
  int kb_per_split;
  const float* bias; int epi;
  int x_vec;                                  // 16-byte cp.async allowed (alignment of X and ldx)
  int d_vec;                                  // 16-byte stores allowed
};

}  // namespace xw
}  // namespace gcnb
