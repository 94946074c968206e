/*
 * TEST INFRASTRUCTURE ONLY — CPU restatement ("oracle") of the reference's GCN
 * aggregation arithmetic.  Nothing in gcn_b200/ may link, load or call this file; only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
 *
 * Parity status: PINNED.  The reference ships no golden vectors, so this
 * restatement is pinned against outputs of the reference itself executed in the build
 * container: torch.spmm on CPU (pygcn/gcn1.py:53) and scipy CSR @ in fp64
 * (tests/golden/make_golden.py writes the fixtures; tests/test_oracle.py checks them).
 *
 * Each function cites the reference code whose arithmetic it restates.
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>

/* C[m x k] = A[m x n] (CSR) * B[n x k], fp32 in / fp32 accumulate, one output element
 * accumulated serially over the row's non-zeros in ascending column order, plain
 * multiply-then-add (no fused multiply-add: oracle/Makefile builds with -ffp-contract=off).
 * Restates the per-row inner loop of the reference kernels,
 *     res += cv.y * bv;                                   flexspmm.cu:72-81 (v31), :474-484 (v35)
 * and the operation the layer asks for, output = torch.spmm(adj, support)
 *                                                         pygcn/gcn1.py:53, pygcn/gcn5.py:90
 */
void oracle_spmm_csr_f32(const int32_t* rowptr, const int32_t* col, const float* val,
                         int m, int n, int k, const float* B, float* C) {
  (void)n;
#pragma omp parallel for schedule(dynamic, 64)
  for (int i = 0; i < m; ++i) {
    float* c = C + (size_t)i * k;
    for (int j = 0; j < k; ++j) c[j] = 0.0f;
    for (int64_t z = rowptr[i]; z < rowptr[i + 1]; ++z) {
      const float a = val[z];
      const float* b = B + (size_t)col[z] * k;
      for (int j = 0; j < k; ++j) c[j] += a * b[j];
    }
  }
}

/* Same product with fp64 accumulation: the "exact" value used to bound the fp32 error of
 * both the reference (whose atomics make its rounding order non-deterministic,
 * flexspmm.cu:85-89) and the CUDA path. */
void oracle_spmm_csr_f64acc(const int32_t* rowptr, const int32_t* col, const float* val,
                            int m, int n, int k, const float* B, double* C) {
  (void)n;
#pragma omp parallel for schedule(dynamic, 64)
  for (int i = 0; i < m; ++i) {
    double* c = C + (size_t)i * k;
    for (int j = 0; j < k; ++j) c[j] = 0.0;
    for (int64_t z = rowptr[i]; z < rowptr[i + 1]; ++z) {
      const double a = (double)val[z];
      const float* b = B + (size_t)col[z] * k;
      for (int j = 0; j < k; ++j) c[j] += a * (double)b[j];
    }
  }
}

/* Transposed product C[n x k] = A^T * B[m x k] (the true backward of torch.spmm; the
 * reference's flexspmm.backward, pygcn/gcn6.py:50-62, reuses A and is only right for a
 * symmetric A — SURVEY.md §4 defect 3).  fp64 accumulate, serial scatter. */
void oracle_spmm_csr_t_f64acc(const int32_t* rowptr, const int32_t* col, const float* val,
                              int m, int n, int k, const float* B, double* C) {
  memset(C, 0, (size_t)n * k * sizeof(double));
  for (int i = 0; i < m; ++i) {
    const float* b = B + (size_t)i * k;
    for (int64_t z = rowptr[i]; z < rowptr[i + 1]; ++z) {
      double* c = C + (size_t)col[z] * k;
      const double a = (double)val[z];
      for (int j = 0; j < k; ++j) c[j] += a * (double)b[j];
    }
  }
}

/* The reference kernel evaluated over its own tile-segment format, sequentially.
 * Restates flexspmm.cu:40-93 (v31; v32-v35 differ only in lane mapping/vector width):
 * for every segment, for each of its tm rows, sum (col,val) pairs stored interleaved in
 * segNzCV with the column id held as a float (tile.cu:67, flexspmm.cu:76), and either
 * store or accumulate into C depending on the MSB of segVoMap (flexspmm.cu:65,85-89).
 * C must be zeroed by the caller, as the reference does (pygcn/gcn6.py:37).            */
void oracle_flexspmm_tiles(const int32_t* seg_rowPtr, const float* segNzCV, const int32_t* segVoMap,
                           int m, int n, int k, int n_segs, int tm, const float* B, float* C) {
  (void)n;
  for (int s = 0; s < n_segs; ++s) {
    for (int r = 0; r < tm; ++r) {
      const int32_t beg = seg_rowPtr[s * (tm + 1) + r];
      const int32_t cnt = seg_rowPtr[s * (tm + 1) + r + 1] - beg;
      const uint32_t tag = (uint32_t)segVoMap[s * tm + r];
      const int atomic = (int)(tag >> 31);
      const int64_t row = (int64_t)(tag & 0x7fffffffu);
      if (row >= m) continue; /* padding rows of the last panel, tile.cu:93-96 */
      for (int j = 0; j < k; ++j) {
        float res = 0.0f;
        for (int z = 0; z < cnt; ++z) {
          const int c = (int)segNzCV[2 * (size_t)(beg + z)];
          const float v = segNzCV[2 * (size_t)(beg + z) + 1];
          res += v * B[(size_t)c * k + j];
        }
        if (atomic) C[row * k + j] += res;
        else        C[row * k + j] = res;
      }
    }
  }
}

/* Row gather X_out[r,:] = X_in[voMp[r],:].  Restates permutate.cu:3-21
 * (flexspmm_v9_permuteX) followed by put_back (:22-39), i.e. the net effect of
 * permutate() (:40-59) written out of place.  Labels are NOT permuted, as in the
 * reference where that branch is disabled (`if (false && ...)`, permutate.cu:17,35). */
void oracle_permute_rows(const float* X_in, const int32_t* voMp, int n, int k, float* X_out) {
  for (int r = 0; r < n; ++r)
    memcpy(X_out + (size_t)r * k, X_in + (size_t)voMp[r] * k, (size_t)k * sizeof(float));
}
