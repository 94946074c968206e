// TEST / BENCH INFRASTRUCTURE ONLY (comparator leg of bench.py; never on the product path).
// cuSPARSE SpMM exactly as the reference calls it — CSR (int32 indices, fp32 values) x row-major dense,
// CUSPARSE_SPMM_CSR_ALG3, alpha = 1, beta = 0 (/root/reference/cuspmm.cu:23-68) — with the per-call
// setup the reference repeats on every call (handle, descriptors, work buffer) hoisted into
// cmp_cusparse_create, so that the timed region of the comparator is the library's SpMM alone.
//   alg: 0 = CUSPARSE_SPMM_CSR_ALG3 (the reference's), 1 = ALG2, 2 = ALG1, 3 = ALG_DEFAULT
#include <cuda_runtime.h>
#include <cusparse.h>
#include <stdlib.h>

struct CmpCusparse {
  cusparseHandle_t handle = nullptr;
  cusparseSpMatDescr_t A = nullptr;
  cusparseDnMatDescr_t B = nullptr, C = nullptr;
  void* buffer = nullptr;
  size_t buffer_bytes = 0;
  cusparseSpMMAlg_t alg = CUSPARSE_SPMM_CSR_ALG3;
  float alpha = 1.f, beta = 0.f;
};

extern "C" {

int cmp_cusparse_create(const int* rowptr_dev, const int* col_dev, const float* val_dev, long long m, long long n,
                        long long nnz, const float* B_dev, float* C_dev, int k, long long ldb, long long ldc, int alg,
                        void* stream, CmpCusparse** out) {
  CmpCusparse* h = new CmpCusparse();
  static const cusparseSpMMAlg_t algs[4] = {CUSPARSE_SPMM_CSR_ALG3, CUSPARSE_SPMM_CSR_ALG2, CUSPARSE_SPMM_CSR_ALG1,
                                            CUSPARSE_SPMM_ALG_DEFAULT};
  h->alg = algs[alg & 3];
  if (cusparseCreate(&h->handle) != CUSPARSE_STATUS_SUCCESS) return 1;
  if (cusparseSetStream(h->handle, (cudaStream_t)stream) != CUSPARSE_STATUS_SUCCESS) return 2;
  if (cusparseCreateCsr(&h->A, m, n, nnz, (void*)rowptr_dev, (void*)col_dev, (void*)val_dev, CUSPARSE_INDEX_32I,
                        CUSPARSE_INDEX_32I, CUSPARSE_INDEX_BASE_ZERO, CUDA_R_32F) != CUSPARSE_STATUS_SUCCESS) return 3;
  if (cusparseCreateDnMat(&h->B, n, k, ldb, (void*)B_dev, CUDA_R_32F, CUSPARSE_ORDER_ROW) != CUSPARSE_STATUS_SUCCESS) return 4;
  if (cusparseCreateDnMat(&h->C, m, k, ldc, (void*)C_dev, CUDA_R_32F, CUSPARSE_ORDER_ROW) != CUSPARSE_STATUS_SUCCESS) return 5;
  if (cusparseSpMM_bufferSize(h->handle, CUSPARSE_OPERATION_NON_TRANSPOSE, CUSPARSE_OPERATION_NON_TRANSPOSE, &h->alpha,
                              h->A, h->B, &h->beta, h->C, CUDA_R_32F, h->alg, &h->buffer_bytes) != CUSPARSE_STATUS_SUCCESS)
    return 6;
  if (h->buffer_bytes && cudaMalloc(&h->buffer, h->buffer_bytes) != cudaSuccess) return 7;
  *out = h;
  return 0;
}

int cmp_cusparse_spmm(CmpCusparse* h) {
  return (int)cusparseSpMM(h->handle, CUSPARSE_OPERATION_NON_TRANSPOSE, CUSPARSE_OPERATION_NON_TRANSPOSE, &h->alpha, h->A,
                           h->B, &h->beta, h->C, CUDA_R_32F, h->alg, h->buffer);
}

long long cmp_cusparse_buffer_bytes(const CmpCusparse* h) { return (long long)h->buffer_bytes; }

void cmp_cusparse_destroy(CmpCusparse* h) {
  if (!h) return;
  cusparseDestroySpMat(h->A); cusparseDestroyDnMat(h->B); cusparseDestroyDnMat(h->C);
  cusparseDestroy(h->handle);
  cudaFree(h->buffer);
  delete h;
}

}  // extern "C"
