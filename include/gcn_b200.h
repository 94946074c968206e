/*
 * gcn_b200 — C ABI of the B200-native GCN aggregation path (libgcn_b200.so).
 *
 * This is the drop-in boundary for the reference's ctypes FFI (pygcn/gcn6.py:21-25 loads five
 * .so files and calls `extern "C" void f(raw pointers, ints)`).  Two layers are exported:
 *
 *   1. gcnb_*  — the native interface: int status codes (0 = ok, <0 = error, text via
 *      gcnb_last_error()), explicit CUDA stream, opaque plan handles, caller-owned buffers.
 *   2. the reference's own symbol names and argument orders (flexspmm, csr2tile,
 *      csr2seg_Cmajor, dfs, gorder, perm_apply, permutate, cuspmm) as a compatibility shim, so
 *      that `ctypes.cdll.LoadLibrary` in the reference's Python can be pointed at this library
 *      unchanged.
 *
 * Plain C types only: no torch, no C++ in any signature.  "file:line" citations are relative
 * to the reference repository (guohaoqiang/gcn).
 */
#ifndef GCN_B200_H
#define GCN_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GCNB_ABI_VERSION 1

/* status codes */
#define GCNB_OK 0
#define GCNB_ERR_ARG (-1)     /* bad argument (null pointer, negative size, unsorted CSR ...) */
#define GCNB_ERR_CUDA (-2)    /* a CUDA runtime/driver call failed; see gcnb_last_error() */
#define GCNB_ERR_NOGPU (-3)   /* no usable sm_100 device: there is NO CPU fallback */
#define GCNB_ERR_ALLOC (-4)
#define GCNB_ERR_UNSUPPORTED (-5)

/* dtype of the dense operand B (features).  Accumulation and C are always fp32. */
#define GCNB_F32 0
#define GCNB_BF16 1

/* epilogue flags for gcnb_spmm / gcnb_xw (fuse pygcn/gcn6.py:142 `output + bias` and
 * :245 F.relu into the producing kernel) */
#define GCNB_EPI_NONE 0
#define GCNB_EPI_BIAS 1
#define GCNB_EPI_RELU 2
#define GCNB_EPI_ACCUM 4 /* gcnb_spmm only: C = C + A·B (then bias/ReLU); used by the column-split
                            multi-GPU path, gcn_b200/dist.py */
#define GCNB_EPI_DROPOUT 8 /* gcnb_xw_ex only: after bias/ReLU, out = keep ? out / (1-p) : 0, keep bits from a counter hash
                              of (seed, row, col / 4), one byte per element, p quantised to 1/256  (F.dropout, gcn6.py:246) */
#define GCNB_EPI_MASK 16   /* gcnb_xw_ex only: out = bit(mask_in, row, col) ? out / (1-p) : 0 — the backward of ReLU followed
                              by dropout, with the bits the forward product wrote (autograd of gcn6.py:245-246) */

/* plan flags */
#define GCNB_PLAN_DEFAULT 0
#define GCNB_PLAN_ASSUME_SYMMETRIC 1 /* skip the check; A^T·dY runs on A (reference behaviour, gcn6.py:50-62) */
#define GCNB_PLAN_NO_TRANSPOSE 2     /* never build A^T (forward-only plans) */
#define GCNB_PLAN_PANEL 4            /* always build the panel class (32-row panels walking their columns in step) */
#define GCNB_PLAN_NO_PANEL 8         /* never build it (the default unless GCNB_PANEL is set in the environment) */

typedef struct gcnb_plan gcnb_plan_t;

int gcnb_abi_version(void);
/* thread-local text of the last error returned on this thread ("" if none) */
const char* gcnb_last_error(void);
/* number of usable CUDA devices; 0 means every compute entry point returns GCNB_ERR_NOGPU */
int gcnb_device_count(void);

/* ---------------------------------------------------------------------------------------------
 * SpMM plan.  Replaces the per-fit tile upload of the reference (pygcn/gcn6.py:334-365: csr2tile
 * on the host, then .to(device) of seg_rowPtr/segNzCV/segVoMap/grouped_tailSeg/next_seg).
 * Input: CSR of A [m x n] in HOST memory, int32 indices, fp32 values, columns ascending within a
 * row (what adj.to_sparse_csr() gives, gcn6.py:302-312).  The plan owns device copies of the
 * matrix in its internal layout and the row-class work lists; it holds no mutable per-launch
 * state, so concurrent launches on different streams are allowed (the reference needs a fresh
 * next_seg copy per launch, gcn6.py:38,53).
 * ------------------------------------------------------------------------------------------- */
int gcnb_plan_create(const int32_t* rowptr, const int32_t* col, const float* val,
                     int64_t m, int64_t n, int64_t nnz, int flags, int device,
                     gcnb_plan_t** plan_out);
int gcnb_plan_destroy(gcnb_plan_t* plan);
/* info[0]=m [1]=n [2]=nnz [3]=symmetric(0/1) [4]=has_transpose [5]=rows in warp class
 * [6]=rows in block class [7]=block work items [8]=max row nnz [9]=device bytes held
 * [10]=SM count [11]=rows split across >1 block item [12]=stream-class segments
 * [13]=rows cut into several stream segments (they add one reduce launch) [14]=panel-class work
 * items (0: class not in use) [15]=distinct columns summed over the 32-row panels (nnz / this = how
 * many rows of a panel share a gathered feature row on average; 0 when the panels were not built) */
int gcnb_plan_info(const gcnb_plan_t* plan, int64_t info[16]);

/* C[m x k] = A · B[n x k]  (+ bias[k]) (ReLU).   Replaces flexspmm() (flexspmm.cu:499-544) and
 * torch.spmm(adj, support) (pygcn/gcn5.py:90, gcn1.py:53).  B, C, bias are DEVICE pointers; B
 * row-major with leading dimension ldb (elements), C row-major with ldc.  C need NOT be zeroed
 * (the reference requires torch.zeros, gcn6.py:37).  No atomics: results are bit-reproducible
 * run to run.  stream = cudaStream_t (torch.cuda.current_stream().cuda_stream), 0 = legacy. */
int gcnb_spmm(const gcnb_plan_t* plan, const void* B, int b_dtype, int64_t ldb,
              float* C, int64_t ldc, int k, const float* bias, int epilogue, void* stream);
/* C[n x k] = A^T · B[m x k]: the true backward of torch.spmm.  Runs on A itself when the plan is
 * symmetric (the reference always does, pygcn/gcn6.py:50-62), else on the transposed copy. */
int gcnb_spmm_t(const gcnb_plan_t* plan, const void* B, int b_dtype, int64_t ldb,
                float* C, int64_t ldc, int k, void* stream);
/* End-to-end form with HOST buffers: H2D of B, SpMM, D2H of C, all on `stream`, synchronised
 * before returning.  transpose != 0 selects A^T.  Device staging is plan-owned and reused. */
int gcnb_spmm_host(gcnb_plan_t* plan, const float* B_host, float* C_host, int k,
                   int transpose, void* stream);

/* out[r,:] = X[vomp[r],:]  — replaces permutate() (permutate.cu:40-59), out of place, no device
 * sync, no malloc.  Labels are not touched (the reference's label branch is disabled,
 * permutate.cu:17,35).  Device pointers. */
int gcnb_permute_rows(const float* X, const int32_t* vomp, float* out, int64_t n, int k,
                      void* stream);

/* Dense feature transform on tcgen05 tensor cores:
 *   D[M x N] = op(X) · op(W) (+bias[N]) (ReLU),  op(X) is M x K, op(W) is K x N.
 * Replaces torch.mm(input, self.weight) (pygcn/gcn6.py:131, :191; gcn5.py:88, :136) and the two
 * products autograd derives from it (dX = dS·W^T: trans_w = 1;  dW = X^T·dS: trans_x = 1).
 * fp32 in / fp32 out, row-major DEVICE pointers with leading dimensions (elements); a set trans
 * flag means the stored matrix is the transpose of the operand (X stored K x M, W stored N x K).
 * `precision` 0 = 3xTF32 split (fp32-parity mode, ~1e-6 rel), 1 = single-pass TF32,
 * 2 = bf16 operands. */
int gcnb_xw(const float* X, int64_t ldx, int trans_x, const float* W, int64_t ldw, int trans_w,
            float* D, int64_t ldd, int64_t M, int N, int K, const float* bias, int epilogue,
            int precision, void* stream);
/* gcnb_xw with the fused elementwise neighbours of the product in the GCN layer (same operands, plus):
 *   GCNB_EPI_DROPOUT: dropout probability `drop_p` in [0,1), 32-bit `seed` (a new one per training step);
 *   mask_out (optional, any epilogue): bit (row, col) of word mask_out[row * ldm + col / 32] = (out > 0) — for a
 *                     ReLU / dropout forward this is the mask its backward needs, 1 bit per element;
 *   GCNB_EPI_MASK:    mask_in in that layout and `drop_p` (scale 1/(1-p); 0 for a plain ReLU mask).
 * gcnb_dropout_keeps returns the keep decision of one element (host function; tests restate the mask with it). */
int gcnb_xw_ex(const float* X, int64_t ldx, int trans_x, const float* W, int64_t ldw, int trans_w,
               float* D, int64_t ldd, int64_t M, int N, int K, const float* bias, int epilogue,
               const uint32_t* mask_in, uint32_t* mask_out, int64_t ldm, float drop_p, uint32_t seed,
               int precision, void* stream);
/* dW = X^T·G [M x N] and db = 1^T·G [N] (the column sums of G) from ONE pass over the vertex-sized operands
 * X [K x M] and G [K x N], both row-major over the K vertices: what autograd derives for `self.weight` and
 * `self.bias` of `torch.mm(input, self.weight)` ... `output + self.bias` (pygcn/gcn6.py:131-143).  On the tall-K
 * streaming kernel the sums ride along as one more row of the product (a ones column in X's tile); elsewhere
 * they are two fixed-order passes over G.  Deterministic either way. */
int gcnb_xw_tn_colsum(const float* X, int64_t ldx, const float* G, int64_t ldg, float* D, int64_t ldd,
                      float* colsum, int64_t M, int N, int K, int precision, void* stream);
int gcnb_dropout_keeps(int64_t row, int col, int N, float drop_p, uint32_t seed);
/* colsum[N] = 1^T·G for a row-major G [K x N], like the column-sum half of gcnb_xw_tn_colsum, but ALLOCATING NOTHING:
 * no cudaMalloc / cudaMallocAsync / cudaFree, no library workspace.  It has no counterpart in the reference; it was
 * added for callers that cannot allocate (code under CUDA-graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc).  The per-slab partial sums go to a static device array of
 * gcnb_colsum_noalloc_max_floats() = 2 Mi floats (8 MiB, held by every GPU that loads the library until exit):
 * ceil(K/512)·N must not exceed it, else GCNB_ERR_ARG with the sizes in gcnb_last_error().  Every call shares the
 * array, so calls that overlap on one GPU (different streams, or one stream per thread) race; calls on one stream
 * are ordered.  Deterministic (fixed summation order), same result as gcnb_xw_tn_colsum's two-pass column sums. */
int gcnb_colsum_noalloc(const float* G, int64_t ldg, int64_t K, int N, float* colsum, void* stream);
int64_t gcnb_colsum_noalloc_max_floats(void);

/* Fused log-softmax + NLL of the GCN head (replaces F.log_softmax, pygcn/gcn6.py:248, and
 * F.nll_loss(output[idx_train], labels[idx_train]), gcn6.py:427) and its gradient.  DEVICE pointers.
 *   fwd: logp[r, 0..ncls) = log_softmax(logits[r, 0..ncls)) for all n rows;
 *        loss[0] = -(1/denom) * sum_{i < n_train} logp[idx_train[i], labels[idx_train[i]]]  (fixed order)
 *   bwd: dlogits[r, c] = gloss[0]/denom * (exp(logp[r,c]) - [c == labels[r]]) where is_train[r], else 0;
 *        columns ncls..ncols-1 (padding of the aggregation width) are written as 0.
 * labels / idx_train are int64 (torch.long), is_train one byte per vertex. */
int gcnb_logsoftmax_nll_fwd(const float* logits, int64_t ldl, int64_t n, int ncls, const int64_t* labels,
                            const int64_t* idx_train, int64_t n_train, int64_t denom, float* logp, int64_t ldo,
                            float* loss, void* stream);
int gcnb_logsoftmax_nll_bwd(const float* logp, int64_t ldo, const int64_t* labels, const uint8_t* is_train,
                            const float* gloss, int64_t denom, int64_t n, int ncls, int ncols, float* dlogits,
                            int64_t ldd, void* stream);


/* ---------------------------------------------------------------------------------------------
 * Multi-GPU exchange over NVLink peer memory (the reference is single-GPU).
 * One process per GPU.  Every rank allocates an exchange buffer (gcnb_peer_alloc), sends the
 * 64-byte handle to its peers by any host channel (torch.distributed), and maps theirs
 * (gcnb_peer_open, CUDA IPC with lazily enabled peer access).  gcnb_pull_rows then copies the
 * halo rows a rank needs straight out of the owners' HBM:
 *     out[i, :] = bases[part[i]][row[i], :]          (row_bytes bytes per row)
 * bases: HOST array of n_parts device pointers (own buffer or opened peers; entries never read may
 * be null); part,row: DEVICE int32[n_rows]; out: DEVICE.  The rows are processed starting at index
 * first_row and wrapping around: with the list sorted by owner, rank p passes the index of its first
 * row owned by rank p+1, so that at any moment the ranks read from DIFFERENT owners instead of all
 * draining GPU 0 first (8 GPUs, products-shaped: 1.75 -> see DESIGN.md §6).  Ordering between ranks (a block must be
 * written before a peer pulls it, and pulled before it is overwritten) is the caller's: see
 * gcn_b200/dist.py.
 * ------------------------------------------------------------------------------------------- */
#define GCNB_IPC_HANDLE_BYTES 64
int gcnb_peer_alloc(int device, int64_t bytes, void** ptr_out, unsigned char* handle_out);
int gcnb_peer_free(int device, void* ptr);
int gcnb_peer_open(int device, const unsigned char* handle, void** ptr_out);
int gcnb_peer_close(int device, void* ptr);
int gcnb_pull_rows(const void* const* bases, int n_parts, int64_t src_pitch_bytes,
                   const int32_t* part, const int32_t* row, void* out, int64_t out_pitch_bytes,
                   int64_t n_rows, int64_t row_bytes, int64_t first_row, void* stream);
/* Stream-ordered rendezvous of the ranks through flags in peer memory (replaces a one-element NCCL all-reduce,
 * which costs ~0.11 ms per SpMM on the critical path, profiles/r02_ah_*).  Every rank owns an array of n_parts
 * uint32 words in a gcnb_peer_alloc'ed buffer, zeroed before first use and mapped on the peers; flag_bases is a
 * HOST array of those n_parts device pointers as seen from THIS process.  The call enqueues one small kernel:
 * it publishes `epoch` into this rank's word of every peer's array (release, system scope) and waits until every
 * peer's word in the rank's own array has reached `epoch`.  All ranks pass the same, growing epoch (1, 2, ...).
 * Work enqueued on `stream` before the call is visible to peers that pass the rendezvous; work after it starts
 * when every rank has arrived.  A peer that never arrives makes the kernel trap after ~10 s. */
int gcnb_peer_barrier(const void* const* flag_bases, int n_parts, int rank, uint32_t epoch, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Host graph preprocessing (bit-exact with the reference).  All arrays HOST memory.
 * rank_out[old] = new, the reference's order_* convention (order_rcm.cu:28-31).
 * ------------------------------------------------------------------------------------------- */
/* order_deg(Edgelist&, desc) order_deg.cu:41-45: in+out degree, ties by vertex id */
int gcnb_order_deg(const int32_t* rowptr, const int32_t* col, int64_t m, int64_t nnz, int desc,
                   int64_t* rank_out);
/* order_rcm(Edgelist&, directed=true) order_rcm.cu:15-33 */
int gcnb_order_rcm(const int32_t* rowptr, const int32_t* col, int64_t m, int64_t nnz,
                   int64_t* rank_out);
/* complete_gorder(Edgelist&, window) order_gorder.cu:13-31 (RCM, then Gorder) */
int gcnb_order_gorder(const int32_t* rowptr, const int32_t* col, int64_t m, int64_t nnz,
                      int window, int64_t* rank_out);
/* Apply vomp[new] = old to a CSR: out = P·A·P^T with columns re-sorted per row.  In place, like
 * perm_apply() renumber.cu:233-318. */
int gcnb_perm_apply(int32_t* rowptr, int32_t* col, float* val, const int32_t* vomp, int64_t m,
                    int64_t nnz);
/* gorder() renumber.cu:157-230 (window 3) and dfs() renumber.cu:23-155: reorder the CSR in place
 * and write vomp[new] = old. */
int gcnb_renumber_gorder(int32_t* rowptr, int32_t* col, float* val, int32_t* vomp, int64_t m,
                         int64_t nnz, int window);
int gcnb_renumber_dfs(int32_t* rowptr, int32_t* col, float* val, int32_t* vomp, int64_t m,
                      int64_t nnz);
/* rabbit() renumber.cu:319-522 — the reference's default order (RBT = 1, pygcn/gcn6.py:27-30):
 * iterative modularity agglomeration (opt_iterative, shyness 1, hub grouping off), vertices laid out
 * by dendrogram leaf order; CSR rewritten in place, vomp[new] = old. */
int gcnb_renumber_rabbit(int32_t* rowptr, int32_t* col, float* val, int32_t* vomp, int64_t m,
                         int64_t nnz);
int gcnb_renumber_rcm(int32_t* rowptr, int32_t* col, float* val, int32_t* vomp, int64_t m,
                      int64_t nnz);
int gcnb_renumber_deg(int32_t* rowptr, int32_t* col, float* val, int32_t* vomp, int64_t m,
                      int64_t nnz, int desc);

/* Reference tile format (tile.cu:11-169), produced for parity and for the compat shim only; the
 * engine's own layout is the plan.  Two-call protocol replaces the reference's guessed buffer
 * sizes (gcn6.py:334-339): gcnb_tile_count returns n_segs, then the caller sizes
 *   seg_rowPtr int32[(tm+1)*n_segs], segNzCV fp32[2*nnz], segVoMap int32[tm*n_segs],
 *   grouped_tailSeg / next_seg int32[n_sm+1].
 * seg_rowPtr is the CUMULATIVE form the kernels index with (flexspmm.cu:63-64); the shipped
 * tile.cu:71 writes a non-cumulative value, see DESIGN.md "tile.cu:71".  n_sm replaces the
 * cudaGetDeviceProperties query of tile.cu:118-122. */
int gcnb_tile_count(const int32_t* rowptr, const int32_t* col, int64_t m, int64_t nnz, int tm,
                    int64_t* n_segs_out);
int gcnb_csr2tile(const int32_t* rowptr, const int32_t* col, const float* val, int64_t m,
                  int64_t n, int64_t nnz, int tm, int n_sm, int32_t* segVoMap,
                  int32_t* seg_rowPtr, float* segNzCV, int32_t* grouped_tailSeg,
                  int32_t* next_seg, int64_t* n_segs_out);

/* ---------------------------------------------------------------------------------------------
 * Compatibility shim: the reference's exported names, argument orders and `void` returns.
 * Errors are recorded for gcnb_last_error() and printed to stderr; nothing aborts.
 * ------------------------------------------------------------------------------------------- */
/* flexspmm.cu:499-502 — device pointers; consumes reference-format tiles; C pre-zeroed by caller */
void flexspmm(int* seg_rowPtr, float* segNzCV, int* segVoMap, int* grouped_tailSeg,
              int* next_seg, int m, int n, int k, int n_segs, float* B, float* C);
/* cuspmm.cu:23-24 — device CSR (rowPtr really int32, passed through the reference's float*) */
void cuspmm(float* rowPtr, int* col, float* vals, float* X, float* C, int m, int n, int nnz,
            int dim);
/* tile.cu:104-106 and :11-12 — host pointers */
void csr2tile(int* rowPtr, int* col, float* vals, int m, int n, int nnz, int* vo_mp,
              int* segVoMap, int* seg_rowPtr, float* segNzCV, int* grouped_tailSeg,
              int* next_seg, int tm, int* n_segs);
void csr2seg_Cmajor(int ridx, int* rowPtr, int* colIdx, float* vals, int m, int n, int nnz,
                    int* voMp, int* segVoMap, int* seg_rowPtr, float* segNzCV, int tm,
                    int* n_segs);
/* renumber.cu:23,157,233,319 — host pointers, CSR rewritten in place */
void dfs(int* rowPtr, int* col, float* vals, int* vomp, int m, int n, int nnz);
void rabbit(int* rowPtr, int* col, float* vals, int* vomp, int m, int n, int nnz);
void gorder(int* rowPtr, int* col, float* vals, int* vomp, int m, int n, int nnz);
void perm_apply(int* rowPtr, int* col, float* vals, int* vomp, int m, int n, int nnz);
/* permutate.cu:40-41 — device pointers, in place (through plan-free scratch), labels untouched */
void permutate(float* B, int* voMp, int* labels, int m, int n, int k);

#ifdef __cplusplus
}
#endif
#endif /* GCN_B200_H */
