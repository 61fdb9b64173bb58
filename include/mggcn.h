/*
 * mggcn.h — C-ABI of the B200-native MG-GCN hot path.
 *
 * This header is the drop-in boundary.  The reference (GT-TDAlab/MG-GCN) has no FFI layer:
 * its closest thing to a C-ABI is the stream-level wrapper layer `(cudaStream_t, raw pointers,
 * sizes)` declared at src/cuda_utils.hpp:398-468, plus the cuSPARSE / cuBLAS / NCCL calls made
 * from src/cuda_utils.hpp:27-32 (SpMM), :158-172 (GEMM) and src/dist_matrix.hpp:458-467
 * (staged broadcast).  Every entry point below names the reference interface it replaces.
 *
 * Conventions
 *  - plain pointers and sizes only; `mggcn_stream_t` is a `cudaStream_t` passed as `void*`;
 *  - all device entry points are asynchronous on the given stream, never synchronise, never
 *    allocate, and borrow their pointers only for the duration of the enqueue;
 *  - return 0 on success, a positive `cudaError_t` on a CUDA failure, a negative MGGCN_E_* on a
 *    usage error.  Nothing in here calls exit() (the reference's CHECK_* macros do,
 *    src/mg_gcn.hpp:31-68; that policy lives in the C++ surface above this ABI);
 *  - there is no CPU fallback: device entry points fail with a CUDA error when no GPU is present.
 */
#ifndef MGGCN_H_
#define MGGCN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MGGCN_API __attribute__((visibility("default")))

typedef void *mggcn_stream_t; /* cudaStream_t */

enum {
    MGGCN_OK = 0,
    MGGCN_E_INVALID = -1,     /* bad argument (shape, alignment, null pointer) */
    MGGCN_E_IO = -2,          /* file could not be opened / short read */
    MGGCN_E_FORMAT = -3,      /* file is not PIGO-CSR-v2 / dense blob, or wrong type widths */
    MGGCN_E_NOMEM = -4,
    MGGCN_E_NCCL = -5,
    MGGCN_E_UNSUPPORTED = -6,
    MGGCN_E_STATE = -7        /* call order violated (e.g. backward before forward) */
};

/* SpMM / GEMM fused epilogues.  The reference runs these as separate kernels over the whole
 * activation matrix (leaky_relu_forward src/cuda_utils.cu:26-31, leaky_relu_backward :33-38,
 * broadcast_rows :40-51); here they are applied to the accumulator before its only store. */
enum {
    MGGCN_EPI_NONE = 0,
    MGGCN_EPI_LEAKY_RELU = 1,     /* out = max(x, slope*x) */
    MGGCN_EPI_LEAKY_RELU_BWD = 2  /* out = aux[r,c] > 0 ? x : slope*x  (aux = the layer output Z) */
};
typedef struct {
    int kind;                 /* MGGCN_EPI_* */
    float slope;              /* LeakyReLU slope (0.01 in the reference, src/cuda_utils.hpp:471,485) */
    const float *bias;        /* optional, F entries: added to every row before the activation */
    const float *aux;         /* LEAKY_RELU_BWD only: rows x F, leading dimension ld_aux */
    int64_t ld_aux;
    const uint8_t *keep_mask; /* optional dropout keep-mask, rows x F bytes, leading dim ld_mask:
                               * out = mask ? out * keep_scale : 0.  Not used by the reference (it
                               * has no dropout); off in every parity run. */
    int64_t ld_mask;
    float keep_scale;
} mggcn_epilogue;

MGGCN_API int mggcn_abi_version(void);
/* Thread-local text for the last non-zero return code of this thread ("" if none). */
MGGCN_API const char *mggcn_last_error(void);
/* Number of kernel launches this library has enqueued in this process (all streams). */
MGGCN_API uint64_t mggcn_launch_count(void);

/* ------------------------------------------------------------------------------------------
 * a1. Local CSR SpMM:  C = alpha * A * B + beta * C  (+ optional fused activation)
 *     replaces  matmul(context, csr_matrix, dn_matrix, dn_matrix, ext_buffer, alpha, beta, alg)
 *     = cusparseSpMM(NON_TRANSPOSE, NON_TRANSPOSE, ALG_DEFAULT, row-major)  src/cuda_utils.hpp:27-32
 *
 *  A: rows x cols CSR, rowptr[rows+1] (u32 or u64), colidx[nnz] u32, vals[nnz] f32, all device.
 *  B: cols x F row-major, leading dimension ldb >= F.   C: rows x F row-major, ldc >= F.
 *  B and C must not alias.  C is read only if beta != 0.
 *  epi:  optional (may be NULL) fused epilogue applied to alpha*A*B + beta*C.
 *  plan: optional (may be NULL) work decomposition from mggcn_spmm_plan_create_*; with a plan
 *        long rows are split and rows are issued longest-first, and rowptr may be NULL; results
 *        are deterministic either way (no atomics).
 *  ws:   workspace of mggcn_spmm_plan_workspace_bytes(plan, F) bytes (NULL if that is 0).
 * ------------------------------------------------------------------------------------------ */
typedef struct mggcn_spmm_plan mggcn_spmm_plan_t;

MGGCN_API int mggcn_spmm_csr_f32_i32(mggcn_stream_t stream, int64_t rows, int64_t cols,
                                     const uint32_t *rowptr, const uint32_t *colidx,
                                     const float *vals, const float *B, int64_t ldb, float *C,
                                     int64_t ldc, int F, float alpha, float beta,
                                     const mggcn_epilogue *epi, const mggcn_spmm_plan_t *plan,
                                     void *ws);
/* int64 row pointers: needed once nnz >= 2^31 (the reference cannot represent that,
 * src/main.cpp:43-45, src/mg_gcn.hpp:93-100). */
MGGCN_API int mggcn_spmm_csr_f32_i64(mggcn_stream_t stream, int64_t rows, int64_t cols,
                                     const uint64_t *rowptr, const uint32_t *colidx,
                                     const float *vals, const float *B, int64_t ldb, float *C,
                                     int64_t ldc, int F, float alpha, float beta,
                                     const mggcn_epilogue *epi, const mggcn_spmm_plan_t *plan,
                                     void *ws);

/* a3 (device side). The aggregation over a whole block row of the 1-D partition while the feature
 * blocks of the other ranks are still arriving:  C = alpha * A[rank][:] * H + beta * C.
 *  A: rows x cols CSR with GLOBAL column ids, every row stored in stage order (own block first, then
 *     blocks rank+1, rank+2, ... cyclically; ascending columns inside a block).
 *  gathered: cols x F (leading dimension ld), global row order, filled block by block by copy engines /
 *     peers while the kernel runs; stage_flags[s] (s = 1..P-1, device memory) becomes non-zero after the
 *     block of stage s = block (rank+s)%P is complete.  B_own: this rank's own rows (same ld), complete
 *     at launch; they are copied into their place of `gathered` on `stream` ahead of the kernel (so
 *     `gathered` is written by this call; pass NULL when they are there already).  cuts: P+1 row
 *     cuts of the partition (host memory).  err (device, nullable) is raised to 2 if a flag does
 *     not arrive within ~4 s of SM clock -- the launch never hangs. */
MGGCN_API int mggcn_spmm_gathered_f32_i32(mggcn_stream_t stream, int64_t rows, int64_t cols,
                                          const uint32_t *rowptr, const uint32_t *colidx,
                                          const float *vals, float *gathered,
                                          const float *B_own, int64_t ld, float *C, int64_t ldc, int F,
                                          float alpha, float beta, int P, int rank,
                                          const uint32_t *cuts, const uint32_t *stage_flags, int *err,
                                          const mggcn_epilogue *epi, const mggcn_spmm_plan_t *plan,
                                          void *ws);

/* a2. replaces get_matmul_buffer / cusparseSpMM_bufferSize (src/cuda_utils.hpp:94-102):
 * the plan is built once per matrix (block) from a HOST copy of the row pointers; it owns a
 * small device-side item list.  split_nnz <= 0 selects the default row-split threshold. */
MGGCN_API int mggcn_spmm_plan_create_i32(int64_t rows, const uint32_t *rowptr_host,
                                         int split_nnz, mggcn_spmm_plan_t **out);
MGGCN_API int mggcn_spmm_plan_create_i64(int64_t rows, const uint64_t *rowptr_host,
                                         int split_nnz, mggcn_spmm_plan_t **out);
/* a2, 2-D schedule.  When B (cols x F floats) is far larger than the L2 every gather of the plain plan is a DRAM
 * sector read (products-shape, F = 256: 122 GB of DRAM traffic for 6 GB of operands).  A blocked plan cuts the
 * columns into blocks whose rows of B -- times a feature window chosen at launch -- stay L2-resident; the SpMM
 * entry points then run (feature windows) x (column blocks) launches over the SAME rowptr / colidx / vals arrays,
 * accumulating C between blocks.  Needs a host copy of colidx; rows must hold their columns ascending in
 * key(c) = (c - col_rot) mod cols and inside [0, col_span) (col_rot = 0, col_span = cols: any sorted CSR).
 * F_hint: the widest F this plan will serve.  Block count: mggcn_tune("blocks", k) asks for exactly k,
 * mggcn_tune("blocks_model", 1) lets a DRAM-traffic model choose; by default the plan has ONE block, because on
 * B200 every blocked configuration measured slower than the plain one. */
MGGCN_API int mggcn_spmm_plan_create_blocked_i32(int64_t rows, int64_t cols, const uint32_t *rowptr_host,
                                                 const uint32_t *colidx_host, uint32_t col_rot,
                                                 int64_t col_span, int F_hint, int split_nnz,
                                                 mggcn_spmm_plan_t **out);
/* number of column blocks of a plan (1 for a plain plan) */
MGGCN_API int mggcn_spmm_plan_blocks(const mggcn_spmm_plan_t *plan);
MGGCN_API size_t mggcn_spmm_plan_workspace_bytes(const mggcn_spmm_plan_t *plan, int F);
MGGCN_API void mggcn_spmm_plan_destroy(mggcn_spmm_plan_t *plan);
/* Tuning knobs for experiments ("unroll", "threads", "policy", "split", "slices", "blocks", "blocks_model", "window",
 * "tile_mb", "tma", "narrow", "gemm_tc", "gemm_kgroup", "skip_exchange", "xstreams" = streams a peer block is split
 * over (0 = by the block size), "xpull" = SM-driven peer pulls, "h2d_sm" = CTAs reading pinned inputs instead of a
 * copy-engine transfer, "h2d_chunk_kb" = size of the pieces an input transfer is issued in (0 = by size on multi-GPU
 * engines, so that peer pulls of the running epoch are not queued behind the whole transfer));
 * a negative value only queries.
 * Returns the old value or MGGCN_E_INVALID for an unknown key. */
MGGCN_API int mggcn_tune(const char *key, int value);
/* Thread-local text naming the kernel instance and schedule the last SpMM call of this thread launched, e.g.
 * "spmm_kernel<16,1,4,4,exact> grid=.. cta=64 F=64 x 2 feature window(s) of 64 x 1 column block(s) of 232968". */
MGGCN_API const char *mggcn_last_spmm_config(void);

/* ------------------------------------------------------------------------------------------
 * a12 / N1. Dense transform, row-major:  C = A(op) * B(op) [+ bias] [activation] [+ beta*C]
 *     replaces  matmul(context, dn, dn, dn, alpha, beta, A_T, B_T) = cublasSgemm (FP32 SIMT)
 *     src/cuda_utils.hpp:158-172, with broadcast_rows (bias pre-fill, src/gcn.hpp:117,221)
 *     and leaky_relu_forward (src/gcn.hpp:449,568) folded into the epilogue.
 *  op(A) is M x K, op(B) is K x N, C is M x N; ta/tb select the transposed operand exactly as
 *  the reference's A_T/B_T flags do (A stored K x M when ta, B stored N x K when tb).
 *  epi (nullable): bias (N entries) is added to every row, then the activation is applied.
 * ------------------------------------------------------------------------------------------ */
MGGCN_API int mggcn_gemm_f32(mggcn_stream_t stream, int ta, int tb, int64_t M, int64_t N,
                             int64_t K, float alpha, const float *A, int64_t lda, const float *B,
                             int64_t ldb, float beta, float *C, int64_t ldc,
                             const mggcn_epilogue *epi, void *ws, size_t ws_bytes);
/* Workspace for the split-K form used when ta != 0 (K = number of vertices): partial products are
 * written per K-slice and added in slice order (deterministic).  0 when none is needed. */
MGGCN_API size_t mggcn_gemm_workspace_bytes(int ta, int tb, int64_t M, int64_t N, int64_t K);
/* G_b = 1^T G  (reference: Sgemm with a ones row, src/gcn.hpp:131,237): out[N] = column sums. */
MGGCN_API int mggcn_colsum_f32(mggcn_stream_t stream, int64_t rows, int64_t N, const float *G,
                               int64_t ldg, float *out, void *ws, size_t ws_bytes);
MGGCN_API size_t mggcn_colsum_workspace_bytes(int64_t rows, int64_t N);
/* Column sums for callers that cannot provide a workspace.  (No counterpart in the reference: it has no
 * allocation-free path at all -- its G_b is a cublasSgemm against a ones row on buffers the layer object
 * cudaMalloc'ed in its constructor.)
 * Who needs it: code that runs where device allocation is off the table -- a host application that handed the whole
 * GPU to its own memory manager and gives plug-ins no allocator, a callback or a capture region in which cudaMalloc
 * is illegal, a latency-bound loop that cannot take the device-wide synchronisation cudaMalloc / cudaFree imply.
 * mggcn_colsum_f32 pushes that problem to the caller (`ws`); this variant keeps the partial sums in a fixed
 * array compiled into the library instead, and calls no cudaMalloc, cudaFree or allocating routine, ever.
 *   Limit: N <= 1024 columns (any number of rows); a wider G returns MGGCN_E_INVALID with nothing enqueued.
 *   Cost: the array is 592 x 1024 floats = 2,424,832 bytes of device memory on every GPU that loads this
 *         library, held until the process exits whether or not this entry point is ever called.
 *   Device: the device that owns `stream` must be current.  The first call on a device resolves the array's
 *         device address (cudaGetSymbolAddress); make that first call before any stream capture starts.
 *   Sharing: every call on a GPU uses the same array.  It still holds the previous call's partials -- harmless,
 *         each call overwrites every entry it later reads -- but two calls that overlap on the GPU (two streams,
 *         two host threads) race on it and both results are garbage; the caller must order them.
 * G, ldg, out and the result are exactly those of mggcn_colsum_f32 (same kernels, same summation order). */
MGGCN_API int mggcn_colsum_static_f32(mggcn_stream_t stream, int64_t rows, int64_t N, const float *G,
                                      int64_t ldg, float *out);

/* ------------------------------------------------------------------------------------------
 * a12. Element-wise kernels adjacent to the SpMM (stream-level mirrors of
 *      src/cuda_utils.hpp:398-468 / src/cuda_utils.cu:229-390).
 * ------------------------------------------------------------------------------------------ */
MGGCN_API int mggcn_leaky_relu_forward_f32(mggcn_stream_t stream, const float *in, float *out,
                                           size_t size, float slope);
/* G_out = in > 0 ? G_in : slope * G_in   (src/cuda_utils.cu:33-38) */
MGGCN_API int mggcn_leaky_relu_backward_f32(mggcn_stream_t stream, const float *in,
                                            const float *G_in, float *G_out, size_t size,
                                            float slope);
/* mat[r,:] = row (discard) or += row   (src/cuda_utils.cu:40-51) */
MGGCN_API int mggcn_broadcast_rows_f32(mggcn_stream_t stream, const float *row, float *mat,
                                       size_t rows, size_t m, int discard);
MGGCN_API int mggcn_axpy_f32(mggcn_stream_t stream, size_t n, float alpha, const float *x,
                             float *y);
/* y = alpha * x + beta * y   (src/cuda_utils.cu:81-87; the SGD step of linear::update, src/gcn.hpp:141-144) */
MGGCN_API int mggcn_axpby_f32(mggcn_stream_t stream, size_t n, float alpha, const float *x, float beta,
                              float *y);

/* ------------------------------------------------------------------------------------------
 * N2. Fused row-softmax + cross-entropy + argmax + gradient, in place on the logits
 *     (replaces the 9 kernels + Sgemm + 2 Sasum of src/gcn.hpp:738-766,890-930).
 *  logits: rows x C (ld); labels int32[rows].  On return logits holds
 *  G = (softmax - onehot(y)) * inv_n.  out[0] = sum |log p_y|, out[1] = #correct over these rows
 *  (device doubles; per-CTA partials in ws are added in CTA order, so the result is deterministic).
 * ------------------------------------------------------------------------------------------ */
MGGCN_API int mggcn_softmax_xent_f32(mggcn_stream_t stream, float *logits, int64_t ld,
                                     int64_t rows, int C, const int32_t *labels, float inv_n,
                                     double *out, void *ws, size_t ws_bytes);
MGGCN_API size_t mggcn_softmax_xent_workspace_bytes(int64_t rows);

/* N2. Fused Adam (replaces Saxpy + 2 axpby + 2 aaxpby + 2 adam_final, src/gcn.hpp:253-279):
 *   i < n_decay: g += wd*p   (the reference decays weights only, not biases, :163,270)
 *   m = (1-b1)*g + b1*m; v = (1-b2)*g*g + b2*v; p -= (lr/bc1) * m / (sqrt(v/bc2)+eps)
 *  g is updated in place like the reference does.  One launch covers a packed [weights | biases]
 *  parameter vector. */
MGGCN_API int mggcn_adam_f32(mggcn_stream_t stream, size_t n, float *p, float *g, float *m,
                             float *v, float lr, float beta1, float beta2, float wd, float eps,
                             float bc1, float bc2, size_t n_decay);

/* ------------------------------------------------------------------------------------------
 * a8 / a9. Host pre-processing (bit-exact with the reference's serial forms).
 *  normalize_cols: data[e] /= sum of column indices[e]      src/matrix.hpp:350-362,377-388
 *  normalize_rows: data[e] /= sum of row                      src/matrix.hpp:341-349
 *  transpose:      counting sort, ascending source row        src/matrix.hpp:426-453
 *  block_split:    P x Q blocks, block-local column ids       src/dist_matrix.hpp:215-259
 *  All arrays are HOST memory; rowptr is uint64 here so one entry point covers both widths
 *  (the C++ surface narrows to x_t).  Outputs are caller-allocated.
 * ------------------------------------------------------------------------------------------ */
MGGCN_API int mggcn_csr_normalize(int64_t rows, int64_t cols, const uint64_t *rowptr,
                                  const uint32_t *colidx, float *vals, int by_column);
MGGCN_API int mggcn_csr_transpose(int64_t rows, int64_t cols, const uint64_t *rowptr,
                                  const uint32_t *colidx, const float *vals, uint64_t *t_rowptr,
                                  uint32_t *t_colidx, float *t_vals);
/* Pass 1 (counts): fills blk_rowptr for every block: blk_rowptr[(i*Q + j)] points to an array of
 * (p[i+1]-p[i]+1) uint64.  Pass 2 (fill): writes colidx/vals of every block. */
MGGCN_API int mggcn_csr_block_counts(int64_t rows, const uint64_t *rowptr, const uint32_t *colidx,
                                     int P, const uint32_t *p, int Q, const uint32_t *q,
                                     uint64_t **blk_rowptr);
MGGCN_API int mggcn_csr_block_fill(int64_t rows, const uint64_t *rowptr, const uint32_t *colidx,
                                   const float *vals, int P, const uint32_t *p, int Q,
                                   const uint32_t *q, uint64_t *const *blk_rowptr,
                                   uint32_t *const *blk_colidx, float *const *blk_vals);

/* prep.py's compute_comm (test/data/prep.py:232-272): needed[i*Q + j] = number of distinct columns of
 * block row i (rows p[i]..p[i+1]) that fall into column block j (q[j]..q[j+1]) -- the rows of H_j that
 * part i gathers from, against the q[j+1]-q[j] rows the staged exchange ships.  Host. */
MGGCN_API int mggcn_csr_needed_rows(int64_t rows, int64_t cols, const uint64_t *rowptr,
                                    const uint32_t *colidx, int P, const uint32_t *p, int Q,
                                    const uint32_t *q, uint64_t *needed);

/* ------------------------------------------------------------------------------------------
 * On-disk formats (test/data/prep.py:46-76; consumer src/matrix.hpp:224-234,486-492).
 *  graph.bin: "PIGO-CSR-v2", u8 sizeof(Label), u8 sizeof(Ordinal), n, m, nrows, ncols,
 *             offsets[n+1] (Ordinal), endpoints[m] (Label), weights[m] f32.
 *  dense:     u32 N, u32 M, N*M elements of elem_bytes (4).
 * ------------------------------------------------------------------------------------------ */
typedef struct {
    int64_t n, nnz, nrows, ncols;
    int label_bytes, ordinal_bytes;
    int has_weights;
} mggcn_csr_file_info;
MGGCN_API int mggcn_csr_file_info_read(const char *path, mggcn_csr_file_info *info);
/* rowptr is always widened to uint64 on read. weights may be NULL. */
MGGCN_API int mggcn_csr_file_read(const char *path, uint64_t *rowptr, uint32_t *colidx,
                                  float *vals);
MGGCN_API int mggcn_csr_file_write(const char *path, int64_t n, int64_t ncols,
                                   const uint64_t *rowptr, const uint32_t *colidx,
                                   const float *vals, int ordinal_bytes);
MGGCN_API int mggcn_dense_file_shape(const char *path, int64_t *n, int64_t *m);
MGGCN_API int mggcn_dense_file_read(const char *path, void *data, size_t elem_bytes);
MGGCN_API int mggcn_dense_file_write(const char *path, int64_t n, int64_t m, const void *data,
                                     size_t elem_bytes);

/* ------------------------------------------------------------------------------------------
 * Synthetic graph generator (stands in for test/data/prep.py, which needs the network).
 * Emulates prep.py:101-126: n and F padded to multiples of 8 with isolated zero rows, one
 * self-loop per node, unit weights, symmetric pattern, then a seeded symmetric permutation.
 *   degree_model 0: uniform-random neighbours;  1: power law (Chung-Lu, exponent `gamma`,
 *   max expected degree `max_deg`).
 * Two-call protocol: gen_graph_begin builds the graph and reports nnz, gen_graph_finish copies
 * it out and frees the state.
 * ------------------------------------------------------------------------------------------ */
typedef struct mggcn_gen mggcn_gen_t;
MGGCN_API int mggcn_gen_graph_begin(int64_t n_raw, int64_t target_directed_edges,
                                    int degree_model, double gamma, int64_t max_deg,
                                    uint64_t seed, int permute, mggcn_gen_t **out,
                                    int64_t *n_padded, int64_t *nnz);
MGGCN_API int mggcn_gen_graph_finish(mggcn_gen_t *g, uint64_t *rowptr, uint32_t *colidx,
                                     float *vals);
/* N4. Rows [row_begin, row_end) of the graph the same arguments give mggcn_gen_graph_begin -- entry for entry the
 * same rows (global column ids), built without the rest of the graph: every rank of a 1-D row partition generates
 * its own block row (papers100M-shape: 3.2 B entries never exist in one place).  Finish with
 * mggcn_gen_graph_finish (rowptr has row_end - row_begin + 1 entries and starts at 0). */
MGGCN_API int mggcn_gen_rows_begin(int64_t n_raw, int64_t target_directed_edges, int degree_model,
                                   double gamma, int64_t max_deg, uint64_t seed, int permute,
                                   int64_t row_begin, int64_t row_end, mggcn_gen_t **out,
                                   int64_t *n_padded, int64_t *nnz_rows);
/* features ~ U(-1,1) for raw rows/cols, zero in the padding; labels uniform in [0,C), 0 in padding.
 * `seed`/`permute` must match the graph's so rows follow the same vertex permutation.
 * Counter-based, so the result does not depend on the thread count. */
MGGCN_API int mggcn_gen_features(int64_t n_raw, int64_t n_padded, int64_t f_raw, int64_t f_padded,
                                 uint64_t seed, int permute, float *out);
MGGCN_API int mggcn_gen_labels(int64_t n_raw, int64_t n_padded, int num_classes, uint64_t seed,
                               int permute, int32_t *out);
/* rows [row_begin, row_end) of the two above (out holds row_end - row_begin rows) */
MGGCN_API int mggcn_gen_features_rows(int64_t n_raw, int64_t n_padded, int64_t f_raw,
                                      int64_t f_padded, uint64_t seed, int permute, int64_t row_begin,
                                      int64_t row_end, float *out);
/* feature rows of `count` arbitrary vertex ids (out: count x f_padded) */
MGGCN_API int mggcn_gen_features_ids(int64_t n_raw, int64_t n_padded, int64_t f_raw, int64_t f_padded,
                                     uint64_t seed, int permute, const uint32_t *ids, int64_t count,
                                     float *out);
MGGCN_API int mggcn_gen_labels_rows(int64_t n_raw, int64_t n_padded, int num_classes, uint64_t seed,
                                    int permute, int64_t row_begin, int64_t row_end, int32_t *out);
/* Weight init of the reference: std::default_random_engine(99) re-seeded per matrix,
 * uniform(-g, g), g = gain * sqrt(3 / rows)   (src/matrix.hpp:539-545). */
MGGCN_API int mggcn_init_weights(int64_t rows, int64_t cols, float gain, float *out);

/* ------------------------------------------------------------------------------------------
 * N3. Epoch engine: one handle per rank (one process per GPU, or P handles in one process).
 * Mirrors gcn / dist_gcn<true,...> (src/gcn.hpp:937-1056) and the epoch loop of
 * src/main.cpp:113-133,153-169.  A and its transpose are given as this rank's block rows
 * (P column blocks each, block-local column ids, HOST memory); the engine uploads them.
 * ------------------------------------------------------------------------------------------ */
typedef struct mggcn_engine mggcn_engine_t;

typedef struct {
    int world;              /* P */
    int rank;               /* this rank (row block) */
    int device;             /* CUDA device ordinal */
    int num_layers;         /* = len(sizes) - 1 */
    const int64_t *sizes;   /* [F_in, d_1..d_L, C]; caller pads C to a multiple of P (src/main.cpp:135) */
    int64_t n_global;       /* N */
    const uint32_t *p;      /* row cuts, P+1 entries (src/main.cpp:139-141) */
    int overlap;            /* 0 = -S (comm on the compute stream), 1 = comm stream */
    int exact_grads;        /* 0 = reference-compatible quirks Q1/Q2, 1 = exact */
    int use_graph;          /* capture the epoch into CUDA graphs */
    const void *nccl_id;    /* 128-byte ncclUniqueId when world > 1 and comm_backend == NCCL */
    float lrelu_slope;      /* 0.01 in the reference */
    int comm_backend;       /* MGGCN_COMM_* */
    int timers;             /* record the reference's named timers (eager mode only) */
} mggcn_engine_config;

/* Exchange back-ends for the staged feature-block transfer (a5) and the gradient reduction:
 *  NCCL: root-rotating ncclBroadcast per stage, exactly the reference's schedule
 *        (src/dist_matrix.hpp:458-467), and one ncclAllReduce for the packed gradients;
 *  P2P:  every rank pulls block (rank+s)%P at stage s out of the owner's arena through
 *        peer-mapped memory (NVLink copy engines) into its place in an all-gather buffer and publishes
 *        a stage flag behind it; ONE SpMM launch per aggregation works through this rank's whole block
 *        row in stage order and waits (on the device) only for blocks that have not landed yet, so
 *        transfer and multiplication overlap without per-stage launches.  Flag barriers in peer
 *        memory order the ranks; gradients are summed in rank order from the peers' arenas.  Needs
 *        mggcn_engine_comm_handle / mggcn_engine_comm_connect after create.
 *  P2P_STAGED: the same transfers with the reference's structure -- one SpMM launch per stage on a
 *        double-buffered staging area (src/cuda_utils.hpp:57-92).
 *  P2P_GROUPED: the transfers of P2P (all-gather buffer), multiplied by plain launches over groups of
 *        stages -- own block | stage 1 | stages 2-3 | 4-7 | 8-15 -- each ordered behind the transfer of
 *        its last block by a stream event: later launches see longer rows and bigger grids (which is what
 *        small blocks lack) while the first ones start as early as the staged form does. */
enum { MGGCN_COMM_NCCL = 0, MGGCN_COMM_P2P = 1, MGGCN_COMM_P2P_STAGED = 2, MGGCN_COMM_P2P_GROUPED = 3 };

MGGCN_API int mggcn_nccl_unique_id(void *id128);
MGGCN_API int mggcn_engine_create(const mggcn_engine_config *cfg, mggcn_engine_t **out);
/* P2P back-end wiring: every rank exports an opaque handle of *len bytes (call with out == NULL to
 * get the length), the host gathers the P handles in rank order (torch.distributed all_gather, or
 * plain memory when all engines live in one process) and hands the concatenation to connect. */
MGGCN_API int mggcn_engine_comm_handle(mggcn_engine_t *e, void *out, size_t cap, size_t *len);
MGGCN_API int mggcn_engine_comm_connect(mggcn_engine_t *e, const void *handles, size_t len);
/* which: 0 = forward operand (blocks of A_T in main.cpp's naming), 1 = backward operand (A).
 * blk_* arrays have P entries: column block j of this rank's block row. */
MGGCN_API int mggcn_engine_set_adjacency(mggcn_engine_t *e, int which,
                                         const uint64_t *const *blk_rowptr,
                                         const uint32_t *const *blk_colidx,
                                         const float *const *blk_vals);
/* Host -> device copy of this rank's rows of X (n_local x F_in) and Y (n_local).
 *  async = 0: copy on the compute stream, return when done;
 *  async = 1: enqueue on the compute stream (ordered after the epochs already queued), return at once;
 *  async = 2: prefetch -- copy into shadow buffers on the engine's copy stream, concurrent with the
 *             epoch the compute stream is running; the next epoch / train_forward swaps them in
 *             (device to device) before it starts.  Pinned host memory expected for 1 and 2. */
MGGCN_API int mggcn_engine_set_inputs(mggcn_engine_t *e, const float *X_host,
                                      const int32_t *Y_host, int async);
MGGCN_API int mggcn_engine_train_forward(mggcn_engine_t *e);
MGGCN_API int mggcn_engine_backward(mggcn_engine_t *e);
MGGCN_API int mggcn_engine_adam_update(mggcn_engine_t *e, float lr, float beta1, float beta2,
                                       float wd, float eps);
/* One whole epoch (forward, loss, backward, Adam) enqueued back to back; no host sync. */
MGGCN_API int mggcn_engine_epoch(mggcn_engine_t *e, float lr, float beta1, float beta2, float wd,
                                 float eps);
/* Captures and instantiates the epoch graph without launching it (no-op when use_graph == 0).
 * A process that drives several engines prepares ALL of them before it launches any: graph
 * instantiation allocates device memory, which waits for peer-mapped GPUs to go idle, while a launched
 * engine spins in its first stage barrier until its peers are launched. */
MGGCN_API int mggcn_engine_prepare(mggcn_engine_t *e, float lr, float beta1, float beta2, float wd,
                                   float eps);
/* Blocks until the rank's streams are idle, then returns this rank's partial sums
 * (sum |log p_y|, number correct) of the last train_forward. */
MGGCN_API int mggcn_engine_loss(mggcn_engine_t *e, double *loss_sum, double *correct);
MGGCN_API int mggcn_engine_sync(mggcn_engine_t *e);
MGGCN_API void *mggcn_engine_stream(mggcn_engine_t *e, int which);
/* Debug / parity access: copies a named tensor to host.  name: "W<l>", "b<l>", "GW<l>", "Gb<l>",
 * "H<l>" (output of layer l, local rows).  Returns the element count in *count. */
MGGCN_API int mggcn_engine_get(mggcn_engine_t *e, const char *name, float *out, size_t cap,
                               size_t *count);
MGGCN_API int mggcn_engine_set(mggcn_engine_t *e, const char *name, const float *in, size_t count);
/* rows [row_begin, row_end) of a named tensor (sampled checks at sizes where a whole activation is gigabytes) */
MGGCN_API int mggcn_engine_get_rows(mggcn_engine_t *e, const char *name, int64_t row_begin, int64_t row_end,
                                    float *out, size_t cap, size_t *count);
/* Timers in the reference's CSV schema "<name>:<ms>\n" (src/matrix.hpp:149-157). */
MGGCN_API int mggcn_engine_dump_timers(mggcn_engine_t *e, const char *prefix, char *buf,
                                       size_t cap, size_t *written);
/* After an eager epoch with timers: one line "<timer name>\t<mggcn_last_spmm_config text>" per SpMM region, i.e.
 * which kernel instance and schedule each aggregation really launched. */
MGGCN_API int mggcn_engine_spmm_configs(mggcn_engine_t *e, char *buf, size_t cap, size_t *written);
/* Switches between graph replay (the product path) and eager execution with named timers. */
MGGCN_API int mggcn_engine_set_mode(mggcn_engine_t *e, int use_graph, int timers);
MGGCN_API int64_t mggcn_engine_launches_per_epoch(mggcn_engine_t *e);
MGGCN_API void mggcn_engine_destroy(mggcn_engine_t *e);

/* a5. Staged broadcast of one feature block (replaces dist_row_dn_matrix::bcast,
 * src/dist_matrix.hpp:458-467): root's `send` (count floats) lands in every rank's `recv`. */
MGGCN_API int mggcn_bcast_stage_f32(void *nccl_comm, mggcn_stream_t stream, const float *send,
                                    float *recv, size_t count, int root);

#ifdef __cplusplus
}
#endif
#endif /* MGGCN_H_ */
