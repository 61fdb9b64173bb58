/*
 * mggcn_oracle.c -- CPU restatement of the MG-GCN hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this file's library.  The product (mg-gcn_b200/, include/) never links or calls it.
 *
 * Every function cites the reference lines (relative to /root/reference) it restates.  It is a
 * restatement, not a copy: plain C, host memory, explicit loops.  Where the reference calls a
 * closed library (cusparseSpMM, cublasSgemm, cublasSasum) the published contract of that call is
 * restated (C = alpha*op(A)*op(B) + beta*C in fp32) with a fixed, documented summation order.
 *
 * Parity pinning: checked against every golden vector the reference's tests hold for this path
 * (test/test_gcn.cpp:98-249 loss/gradient known answers incl. test_csr_g, the only test through
 * cusparseSpMM; test/test_matrix.cpp:64-76,93-109; fixtures test/data/toyA|toyB) in
 * tests/test_oracle_golden.py.  At realistic sizes the reference pins nothing, so
 * large-size parity is "oracle vs CUDA path" here plus "reference binary vs CUDA path" on the GPU
 * box (oracle/_ref, built by oracle/Makefile from the sources where they lie).
 *
 * OpenMP is used over independent rows only; every per-row result is a fixed serial sum, so the
 * output does not depend on the thread count.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

ORC_API int orc_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ---------------------------------------------------------------------------------------------
 * csr_matrix::normalize -- serial form, src/matrix.hpp:364-388 (the par_unseq form at :340-362
 * races on in_degree under TBB; the serial form is the definition).
 * axis == 0: divide each entry by its row sum; axis != 0: by its column sum, both accumulated in
 * fp32 in CSR order.
 * ------------------------------------------------------------------------------------------- */
ORC_API void orc_normalize(int64_t rows, int64_t cols, const uint64_t *rowptr,
                           const uint32_t *colidx, float *vals, int axis) {
    if (!axis) {
        for (int64_t v = 0; v < rows; ++v) {
            float sum = 0;
            for (uint64_t e = rowptr[v]; e < rowptr[v + 1]; ++e) sum += vals[e];
            for (uint64_t e = rowptr[v]; e < rowptr[v + 1]; ++e) vals[e] /= sum;
        }
    } else {
        float *in_degree = (float *)calloc((size_t)cols, sizeof(float));
        for (int64_t v = 0; v < rows; ++v)
            for (uint64_t e = rowptr[v]; e < rowptr[v + 1]; ++e) in_degree[colidx[e]] += vals[e];
        for (int64_t v = 0; v < rows; ++v)
            for (uint64_t e = rowptr[v]; e < rowptr[v + 1]; ++e) vals[e] /= in_degree[colidx[e]];
        free(in_degree);
    }
}

/* ---------------------------------------------------------------------------------------------
 * csr_matrix::transpose -- serial form, src/matrix.hpp:426-453: counting sort by column; inside
 * an output row entries appear in ascending source-row order.
 * ------------------------------------------------------------------------------------------- */
ORC_API void orc_transpose(int64_t rows, int64_t cols, const uint64_t *rowptr,
                           const uint32_t *colidx, const float *vals, uint64_t *t_rowptr,
                           uint32_t *t_colidx, float *t_vals) {
    const uint64_t nnz = rowptr[rows];
    uint64_t *dloc = (uint64_t *)malloc((size_t)(nnz ? nnz : 1) * sizeof(uint64_t));
    for (int64_t c = 0; c <= cols; ++c) t_rowptr[c] = 0;
    for (int64_t i = 0; i < rows; ++i)
        for (uint64_t j = rowptr[i]; j < rowptr[i + 1]; ++j) dloc[j] = t_rowptr[colidx[j] + 1]++;
    for (int64_t c = 0; c < cols; ++c) t_rowptr[c + 1] += t_rowptr[c]; /* inclusive scan */
    for (int64_t i = 0; i < rows; ++i)
        for (uint64_t j = rowptr[i]; j < rowptr[i + 1]; ++j) {
            const uint64_t loc = t_rowptr[colidx[j]] + dloc[j];
            t_colidx[loc] = (uint32_t)i;
            t_vals[loc] = vals[j];
        }
    free(dloc);
}

/* ---------------------------------------------------------------------------------------------
 * dist_row_csr_matrix ctor -- src/dist_matrix.hpp:215-259, one block (i, j) at a time:
 * rows [r0, r1), columns [c0, c1); per-row counts -> inclusive scan -> fill; entries keep their
 * within-row order; stored column = c - c0 (block-local, :250).
 * Call with out_colidx == NULL to get only out_rowptr (r1 - r0 + 1 entries).
 * ------------------------------------------------------------------------------------------- */
ORC_API void orc_block(const uint64_t *rowptr, const uint32_t *colidx, const float *vals,
                       int64_t r0, int64_t r1, uint32_t c0, uint32_t c1, uint64_t *out_rowptr,
                       uint32_t *out_colidx, float *out_vals) {
    out_rowptr[0] = 0;
    for (int64_t r = r0; r < r1; ++r) {
        uint64_t k = 0;
        for (uint64_t e = rowptr[r]; e < rowptr[r + 1]; ++e) k += (colidx[e] >= c0 && colidx[e] < c1);
        out_rowptr[r - r0 + 1] = out_rowptr[r - r0] + k;
    }
    if (!out_colidx) return;
    for (int64_t r = r0; r < r1; ++r) {
        uint64_t o = out_rowptr[r - r0];
        for (uint64_t e = rowptr[r]; e < rowptr[r + 1]; ++e)
            if (colidx[e] >= c0 && colidx[e] < c1) {
                out_colidx[o] = colidx[e] - c0;
                out_vals[o] = vals[e];
                ++o;
            }
    }
}

/* csr_matrix::as_dn -- src/matrix.hpp:328-337 */
ORC_API void orc_csr_to_dense(int64_t rows, int64_t cols, const uint64_t *rowptr,
                              const uint32_t *colidx, const float *vals, float *out) {
    memset(out, 0, (size_t)rows * cols * sizeof(float));
    for (int64_t v = 0; v < rows; ++v)
        for (uint64_t e = rowptr[v]; e < rowptr[v + 1]; ++e) out[v * cols + colidx[e]] = vals[e];
}

/* ---------------------------------------------------------------------------------------------
 * SpMM  C = alpha*A*B + beta*C -- contract of matmul(context, csr, dn, dn, ...) =
 * cusparseSpMM(NON_TRANSPOSE, NON_TRANSPOSE), row-major B and C, src/cuda_utils.hpp:16-32.
 * cuSPARSE's summation order is unspecified; this restatement sums each row's products in CSR
 * order in fp32 (acc64 == 0) or in fp64 (acc64 != 0, used to budget rounding error).
 *
 * nblk > 1 restates the staged multi-GPU form, src/cuda_utils.hpp:57-92: for stage i = 0..P-1 the
 * block partial (entries with cuts[i] <= col < cuts[i+1]) is formed from zero and then added into
 * C (beta only on stage 0, :82-83).
 * ------------------------------------------------------------------------------------------- */
ORC_API void orc_spmm(int64_t rows, const uint64_t *rowptr, const uint32_t *colidx,
                      const float *vals, const float *B, int64_t ldb, float *C, int64_t ldc, int F,
                      float alpha, float beta, int acc64, int nblk, const uint32_t *cuts) {
#pragma omp parallel
    {
        float *acc = (float *)malloc((size_t)F * sizeof(float));
        double *dacc = (double *)malloc((size_t)F * sizeof(double));
#pragma omp for schedule(dynamic, 64)
        for (int64_t r = 0; r < rows; ++r) {
            float *c = C + r * ldc;
            if (nblk <= 1) {
                if (acc64) {
                    for (int f = 0; f < F; ++f) dacc[f] = 0;
                    for (uint64_t e = rowptr[r]; e < rowptr[r + 1]; ++e) {
                        const float *b = B + (int64_t)colidx[e] * ldb;
                        const double v = vals[e];
                        for (int f = 0; f < F; ++f) dacc[f] += v * (double)b[f];
                    }
                    for (int f = 0; f < F; ++f)
                        c[f] = (float)((double)alpha * dacc[f] + (beta != 0.f ? (double)beta * c[f] : 0.0));
                } else {
                    for (int f = 0; f < F; ++f) acc[f] = 0;
                    for (uint64_t e = rowptr[r]; e < rowptr[r + 1]; ++e) {
                        const float *b = B + (int64_t)colidx[e] * ldb;
                        const float v = vals[e];
                        for (int f = 0; f < F; ++f) acc[f] += v * b[f];
                    }
                    for (int f = 0; f < F; ++f) c[f] = alpha * acc[f] + (beta != 0.f ? beta * c[f] : 0.f);
                }
            } else {
                for (int i = 0; i < nblk; ++i) {
                    for (int f = 0; f < F; ++f) acc[f] = 0;
                    for (uint64_t e = rowptr[r]; e < rowptr[r + 1]; ++e) {
                        if (colidx[e] < cuts[i] || colidx[e] >= cuts[i + 1]) continue;
                        const float *b = B + (int64_t)colidx[e] * ldb;
                        const float v = vals[e];
                        for (int f = 0; f < F; ++f) acc[f] += v * b[f];
                    }
                    const float bt = i == 0 ? beta : 1.f;
                    for (int f = 0; f < F; ++f) c[f] = alpha * acc[f] + (bt != 0.f ? bt * c[f] : 0.f);
                }
            }
        }
        free(acc);
        free(dacc);
    }
}

/* ---------------------------------------------------------------------------------------------
 * GEMM  C = alpha*op(A)*op(B) + beta*C, row-major -- contract of matmul(context, dn, dn, dn,
 * alpha, beta, A_T, B_T) = cublasSgemm with swapped operands, src/cuda_utils.hpp:158-172.
 * op(A) is M x K, op(B) is K x N.  fp32 accumulation, k ascending.
 * ------------------------------------------------------------------------------------------- */
ORC_API void orc_gemm(int ta, int tb, int64_t M, int64_t N, int64_t K, float alpha, const float *A,
                      int64_t lda, const float *B, int64_t ldb, float beta, float *C, int64_t ldc) {
    if (!ta) {
#pragma omp parallel
        {
            float *acc = (float *)malloc((size_t)N * sizeof(float));
#pragma omp for schedule(static)
            for (int64_t i = 0; i < M; ++i) {
                for (int64_t j = 0; j < N; ++j) acc[j] = 0;
                if (!tb) {
                    for (int64_t k = 0; k < K; ++k) {
                        const float a = A[i * lda + k];
                        const float *b = B + k * ldb;
                        for (int64_t j = 0; j < N; ++j) acc[j] += a * b[j];
                    }
                } else {
                    for (int64_t j = 0; j < N; ++j) {
                        const float *a = A + i * lda, *b = B + j * ldb;
                        float s = 0;
                        for (int64_t k = 0; k < K; ++k) s += a[k] * b[k];
                        acc[j] = s;
                    }
                }
                float *c = C + i * ldc;
                for (int64_t j = 0; j < N; ++j) c[j] = alpha * acc[j] + (beta != 0.f ? beta * c[j] : 0.f);
            }
            free(acc);
        }
    } else {
        /* op(A) = A^T with A stored K x M (the G_W = X^T G shape: K = rows of X, huge).  The
         * k-ascending order per output element is kept; threads split the M x N outputs. */
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < M; ++i) {
            float *c = C + i * ldc;
            float *acc = (float *)calloc((size_t)N, sizeof(float));
            for (int64_t k = 0; k < K; ++k) {
                const float a = A[k * lda + i];
                if (!tb) {
                    const float *b = B + k * ldb;
                    for (int64_t j = 0; j < N; ++j) acc[j] += a * b[j];
                } else {
                    for (int64_t j = 0; j < N; ++j) acc[j] += a * B[j * ldb + k];
                }
            }
            for (int64_t j = 0; j < N; ++j) c[j] = alpha * acc[j] + (beta != 0.f ? beta * c[j] : 0.f);
            free(acc);
        }
    }
}

/* leaky_relu_forward_kernel src/cuda_utils.cu:26-31 */
ORC_API void orc_leaky_relu_forward(const float *in, float *out, size_t n, float alpha) {
    for (size_t i = 0; i < n; ++i) {
        const float x = in[i], ax = alpha * x;
        out[i] = x > ax ? x : ax;
    }
}
/* leaky_relu_backward_kernel src/cuda_utils.cu:33-38 */
ORC_API void orc_leaky_relu_backward(const float *in, const float *g_in, float *g_out, size_t n,
                                     float alpha) {
    for (size_t i = 0; i < n; ++i) g_out[i] = in[i] > 0 ? g_in[i] : alpha * g_in[i];
}
/* broadcast_rows_kernel src/cuda_utils.cu:40-51 */
ORC_API void orc_broadcast_rows(const float *row, float *mat, size_t rows, size_t m, int discard) {
    for (size_t r = 0; r < rows; ++r)
        for (size_t j = 0; j < m; ++j) {
            if (discard) mat[r * m + j] = row[j];
            else mat[r * m + j] += row[j];
        }
}

/* ---------------------------------------------------------------------------------------------
 * Loss -- softmax_cross_entropy_loss::operator(), src/gcn.hpp:785-818 (single GPU) and
 * dist_row_softmax_cross_entropy_loss, :890-930, with softmax :651-675 / :738-766 and kernels
 * max_rows (cuda_utils.cu:95-104), subtract_rows_exp (:192-200), Sgemm with ones (row sum,
 * ascending column order here), scale_rows (:75-79), max_row_indices (:119-133, first maximum,
 * strict <), index_log_rows (:142-150), add_indexed_rows(-1) (:159-164), Sscal(1/N), is_equal
 * (:180-184), Sasum.  In place on `logits`; on return it holds G = (softmax - onehot)/n_global.
 * Returns sum |log p_y| and the number of correct rows over these `rows` (the caller divides by
 * N, src/gcn.hpp:817,929).  The reference compiles with -use_fast_math (CMakeLists.txt:57), so
 * expf/logf/divide there are the approximate intrinsics: ~2 ulp away from this libm form.
 * ------------------------------------------------------------------------------------------- */
ORC_API void orc_softmax_xent(float *logits, int64_t rows, int C, const int32_t *labels,
                              int64_t n_global, float *loss_sum, float *correct) {
    float ls = 0, cs = 0;
    const float inv_n = (float)1 / (float)n_global;
    for (int64_t r = 0; r < rows; ++r) {
        float *h = logits + r * (int64_t)C;
        float mx = -INFINITY;
        for (int i = 0; i < C; ++i) mx = mx > h[i] ? mx : h[i];
        float s = 0;
        for (int i = 0; i < C; ++i) {
            h[i] = expf(h[i] - mx);
            s += h[i];
        }
        for (int i = 0; i < C; ++i) h[i] /= s;
        float best = -INFINITY;
        int arg = 0;
        for (int i = 0; i < C; ++i)
            if (best < h[i]) {
                best = h[i];
                arg = i;
            }
        ls += fabsf(logf(h[labels[r]]));
        cs += (float)(arg == labels[r]);
        h[labels[r]] += (float)-1;
        for (int i = 0; i < C; ++i) h[i] *= inv_n;
    }
    *loss_sum = ls;
    *correct = cs;
}

/* ---------------------------------------------------------------------------------------------
 * Adam -- linear::adam_update src/gcn.hpp:146-172 / dist_row_linear :253-279 with kernels axpby
 * (cuda_utils.cu:81-86), aaxpby (:88-93), adam_final (:208-218; wrapper :378-383 passes
 * step_size = lr / bc1, c2 = bc2).  wd == 0 for the bias (only G_W gets the Saxpy, :163,270).
 * ------------------------------------------------------------------------------------------- */
ORC_API void orc_adam(size_t n, float *p, float *g, float *m, float *v, float lr, float beta1,
                      float beta2, float wd, float eps, int step) {
    const float bc1 = (float)(1 - pow(beta1, step));
    const float bc2 = (float)(1 - pow(beta2, step));
    const float a1 = 1 - beta1, a2 = 1 - beta2;
    const float step_size = lr / bc1;
    for (size_t i = 0; i < n; ++i) {
        if (wd != 0.f) g[i] += wd * p[i];
        m[i] = a1 * g[i] + beta1 * m[i];
        v[i] = a2 * g[i] * g[i] + beta2 * v[i];
        const float denom = sqrtf(v[i] / bc2) + eps;
        p[i] -= step_size * m[i] / denom;
    }
}

/* ---------------------------------------------------------------------------------------------
 * Weight init -- dn_matrix::init src/matrix.hpp:539-545: std::default_random_engine gen(99)
 * re-seeded per matrix, gain *= sqrt(3.0 / rows), std::uniform_real_distribution<float>(-g, g).
 * libstdc++ (GCC 13, the toolchain the reference builds with here): default_random_engine is
 * minstd_rand0 (x <- 16807 x mod 2^31-1, min 1, max 2^31-2); generate_canonical<float,24> draws
 * once: (float)(x - 1) / (float)2147483646, clamped below 1; the distribution returns
 * canon * (b - a) + a.  First four draws for rows = 4, default gain: -1.22278583 -1.17020714
 * 0.746877909 -0.229351223 -- asserted in tests.
 * ------------------------------------------------------------------------------------------- */
ORC_API void orc_init_weights(int64_t rows, int64_t cols, float gain, float *out) {
    uint64_t x = 99 % 2147483647ull;
    if (x == 0) x = 1;
    gain = (float)((double)gain * sqrt(3.0 / (double)rows));
    const float a = -gain, b = gain;
    const float range = (float)2147483646.0;
    for (int64_t i = 0; i < rows * cols; ++i) {
        x = (16807ull * x) % 2147483647ull;
        float canon = (float)(x - 1) / range;
        if (canon >= 1.0f) canon = nextafterf(1.0f, 0.0f);
        /* uniform_real_distribution: canonical * (b - a) + a, contracted into ONE fma by g++ -O3 on any
         * FMA-capable target (how the reference is built: -march=native; oracle/_ref: -march=x86-64-v3);
         * verified bit for bit against the reference build's W / b dumps (tests/test_ref_layers.py) */
        out[i] = fmaf(canon, b - a, a);
    }
}

/* =============================================================================================
 * The model: gcn (src/gcn.hpp:937-995) and dist_gcn<true,...> (:997-1056) as one serial program
 * over the full matrices.  P > 1 only changes (i) the staged SpMM summation (orc_spmm nblk),
 * (ii) G_W / G_b formed as P per-rank partials summed in rank order (allreduce, :238-240,
 * cuda_utils.hpp:304-313), (iii) the loss partial sums (:920-929).
 * ============================================================================================= */
typedef struct {
    int64_t in, out;
    int activation, backward_spmm, gemm_first;
    float *W, *b, *GW, *Gb, *mW, *vW, *mb, *vb;
    float *AHW;       /* layer output buffer, N x max(in,out): also G_out (gcn.hpp:433-434,552-553) */
    float *AH;        /* exact mode only: saved A*H for SpMM-first layers */
    const float *H;   /* layer input (borrowed) */
    const float *X;   /* GEMM input saved by lin(): H (gemm-first) or HW (spmm-first) */
} orc_layer;

typedef struct {
    int L; /* number of layers */
    int64_t N;
    int P;
    uint32_t *cuts;
    const uint64_t *f_rowptr, *b_rowptr; /* forward operand (A_T in main.cpp) / backward operand (A) */
    const uint32_t *f_colidx, *b_colidx;
    const float *f_vals, *b_vals;
    orc_layer *layers;
    float *HW;    /* shared HW_buffer N x max_d (gcn.hpp:949-952,1013-1018) */
    float *G;     /* gradient flowing backwards (aliases the last AHW like the reference) */
    int exact;
    int step;
    float slope;
} orc_gcn;

ORC_API orc_gcn *orc_gcn_create(int num_layers, const int64_t *sizes, int64_t N, int P,
                                const uint32_t *cuts, const uint64_t *f_rowptr,
                                const uint32_t *f_colidx, const float *f_vals,
                                const uint64_t *b_rowptr, const uint32_t *b_colidx,
                                const float *b_vals, int exact_grads) {
    orc_gcn *g = (orc_gcn *)calloc(1, sizeof(orc_gcn));
    g->L = num_layers;
    g->N = N;
    g->P = P;
    g->cuts = (uint32_t *)malloc((size_t)(P + 1) * sizeof(uint32_t));
    for (int i = 0; i <= P; ++i) g->cuts[i] = cuts ? cuts[i] : (uint32_t)((int64_t)i * N / P);
    g->f_rowptr = f_rowptr; g->f_colidx = f_colidx; g->f_vals = f_vals;
    g->b_rowptr = b_rowptr; g->b_colidx = b_colidx; g->b_vals = b_vals;
    g->exact = exact_grads;
    g->slope = 0.01f; /* leaky_relu default alpha, src/cuda_utils.hpp:471,485 */
    g->layers = (orc_layer *)calloc((size_t)num_layers, sizeof(orc_layer));
    int64_t max_d = 0;
    for (int l = 0; l < num_layers; ++l) {
        int64_t d = sizes[l] < sizes[l + 1] ? sizes[l] : sizes[l + 1];
        if (d > max_d) max_d = d;
    }
    g->HW = (float *)calloc((size_t)(N * max_d), sizeof(float));
    for (int l = 0; l < num_layers; ++l) {
        orc_layer *y = &g->layers[l];
        y->in = sizes[l];
        y->out = sizes[l + 1];
        y->activation = l + 1 < num_layers;   /* i + 1 < sizes.size(), gcn.hpp:954,1023 */
        y->backward_spmm = l != 0;            /* i != 1 in the 1-based loop */
        y->gemm_first = y->out <= y->in;      /* HW.m() == AHW.m(), gcn.hpp:439,558 */
        const size_t wn = (size_t)(y->in * y->out);
        y->W = (float *)malloc(wn * sizeof(float));
        y->b = (float *)malloc((size_t)y->out * sizeof(float));
        orc_init_weights(y->in, y->out, (float)sqrt(2 / (1 + 0.01 * 0.01)), y->W); /* gcn.hpp:108,212 */
        orc_init_weights(1, y->out, sqrtf((float)1.0 / 3), y->b);                  /* gcn.hpp:109,213 */
        y->GW = (float *)calloc(wn, sizeof(float));
        y->Gb = (float *)calloc((size_t)y->out, sizeof(float));
        y->mW = (float *)calloc(wn, sizeof(float));
        y->vW = (float *)calloc(wn, sizeof(float));
        y->mb = (float *)calloc((size_t)y->out, sizeof(float));
        y->vb = (float *)calloc((size_t)y->out, sizeof(float));
        const int64_t w = y->in > y->out ? y->in : y->out;
        y->AHW = (float *)calloc((size_t)(N * w), sizeof(float));
        if (exact_grads && !y->gemm_first) y->AH = (float *)calloc((size_t)(N * y->in), sizeof(float));
    }
    return g;
}

ORC_API void orc_gcn_destroy(orc_gcn *g) {
    if (!g) return;
    for (int l = 0; l < g->L; ++l) {
        orc_layer *y = &g->layers[l];
        free(y->W); free(y->b); free(y->GW); free(y->Gb); free(y->mW); free(y->vW);
        free(y->mb); free(y->vb); free(y->AHW); free(y->AH);
    }
    free(g->layers); free(g->HW); free(g->cuts); free(g);
}

static void orc_fwd_spmm(orc_gcn *g, const float *B, float *C, int F) {
    orc_spmm(g->N, g->f_rowptr, g->f_colidx, g->f_vals, B, F, C, F, F, 1.f, 0.f, 0, g->P, g->cuts);
}
static void orc_bwd_spmm(orc_gcn *g, const float *B, float *C, int F) {
    orc_spmm(g->N, g->b_rowptr, g->b_colidx, g->b_vals, B, F, C, F, F, 1.f, 0.f, 0, g->P, g->cuts);
}

/* lin(ctx, X, XW): broadcast_rows(b -> XW) then Sgemm beta = 1 (gcn.hpp:116-123,220-227) */
static void orc_lin_forward(orc_layer *y, int64_t N, const float *X, float *XW) {
    orc_broadcast_rows(y->b, XW, (size_t)N, (size_t)y->out, 1);
    orc_gemm(0, 0, N, y->out, y->in, 1.f, X, y->in, y->W, y->out, 1.f, XW, y->out);
    y->X = X;
}

/* lin.backward (gcn.hpp:125-139,229-246): G_b = 1^T G, G_W = X^T G, [G_out = G W^T]; for P > 1
 * each rank forms its partial over its rows, then the partials are summed (allreduce). */
static void orc_lin_backward(orc_gcn *g, orc_layer *y, const float *X, const float *G, float *G_out,
                             int backward_out) {
    const int64_t N = g->N, in = y->in, out = y->out;
    float *pw = (float *)malloc((size_t)(in * out) * sizeof(float));
    memset(y->GW, 0, (size_t)(in * out) * sizeof(float));
    memset(y->Gb, 0, (size_t)out * sizeof(float));
    for (int r = 0; r < g->P; ++r) {
        const int64_t r0 = g->cuts[r], r1 = g->cuts[r + 1];
        /* G_b partial: ones(1 x n) * G */
        for (int64_t j = 0; j < out; ++j) {
            float s = 0;
            for (int64_t i = r0; i < r1; ++i) s += G[i * out + j];
            y->Gb[j] = g->P == 1 ? s : y->Gb[j] + s;
        }
        orc_gemm(1, 0, in, out, r1 - r0, 1.f, X + r0 * in, in, G + r0 * out, out, 0.f, pw, out);
        for (int64_t k = 0; k < in * out; ++k) y->GW[k] = g->P == 1 ? pw[k] : y->GW[k] + pw[k];
    }
    free(pw);
    if (backward_out) orc_gemm(0, 1, N, in, out, 1.f, G, out, y->W, out, 0.f, G_out, in);
}

/* gcn::train_forward (gcn.hpp:965-974, 1026-1035); layer forward :437-458 / :556-577.
 * X is N x sizes[0], Y has N labels.  Returns (loss, acc) as the reference prints them. */
ORC_API void orc_gcn_train_forward(orc_gcn *g, const float *X, const int32_t *Y, float *loss,
                                   float *acc) {
    const int64_t N = g->N;
    const float *H = X;
    for (int l = 0; l < g->L; ++l) {
        orc_layer *y = &g->layers[l];
        y->H = H;
        if (y->gemm_first) {
            orc_lin_forward(y, N, H, g->HW);              /* HW = H W + b */
            orc_fwd_spmm(g, g->HW, y->AHW, (int)y->out);  /* AHW = A HW */
        } else {
            float *AH = g->exact ? y->AH : g->HW;
            orc_fwd_spmm(g, H, AH, (int)y->in);           /* HW = A H */
            orc_lin_forward(y, N, AH, y->AHW);            /* AHW = HW W + b */
        }
        if (y->activation) orc_leaky_relu_forward(y->AHW, y->AHW, (size_t)(N * y->out), g->slope);
        H = y->AHW;
    }
    /* loss in place on the last AHW (copy = residual_layer = false, gcn.hpp:946,1012) */
    orc_layer *last = &g->layers[g->L - 1];
    float ls = 0, cs = 0;
    for (int r = 0; r < g->P; ++r) {
        const int64_t r0 = g->cuts[r], r1 = g->cuts[r + 1];
        float l1, c1;
        orc_softmax_xent(last->AHW + r0 * last->out, r1 - r0, (int)last->out, Y + r0, N, &l1, &c1);
        ls += l1;
        cs += c1;
    }
    g->G = last->AHW;
    *loss = ls / (float)N;
    *acc = cs / (float)N;
}

/* gcn::backward (gcn.hpp:976-980,1037-1041); layer backward :460-489 / :579-608, including the
 * reference's gradient quirks: Q1 first layer skips the backward SpMM, Q2
 * SpMM-first layers use the raw input H as the GEMM input.  exact != 0 removes both. */
ORC_API void orc_gcn_backward(orc_gcn *g) {
    const int64_t N = g->N;
    float *G = g->G;
    for (int l = g->L - 1; l >= 0; --l) {
        orc_layer *y = &g->layers[l];
        float *T = G;
        if (y->activation) {
            orc_leaky_relu_backward(y->AHW, G, y->AHW, (size_t)(N * y->out), g->slope);
            T = y->AHW;
        }
        const int first = !y->backward_spmm;
        float *G_out = y->AHW; /* G_out aliases AHW_buffer */
        if (y->gemm_first) {
            const float *G_HW = T;
            if (!first || g->exact) {
                orc_bwd_spmm(g, T, g->HW, (int)y->out);
                G_HW = g->HW;
            }
            orc_lin_backward(g, y, y->H, G_HW, G_out, !first);
        } else {
            const float *Xin = g->exact ? y->AH : y->H; /* lin.setX(H), gcn.hpp:477,596 */
            orc_lin_backward(g, y, Xin, T, g->HW, !first);
            if (!first) orc_bwd_spmm(g, g->HW, G_out, (int)y->in);
        }
        G = G_out;
    }
}

/* gcn::adam_update gcn.hpp:987-990,1048-1051 */
ORC_API void orc_gcn_adam(orc_gcn *g, float lr, float beta1, float beta2, float wd, float eps) {
    g->step += 1;
    for (int l = 0; l < g->L; ++l) {
        orc_layer *y = &g->layers[l];
        orc_adam((size_t)(y->in * y->out), y->W, y->GW, y->mW, y->vW, lr, beta1, beta2, wd, eps, g->step);
        orc_adam((size_t)y->out, y->b, y->Gb, y->mb, y->vb, lr, beta1, beta2, 0.f, eps, g->step);
    }
}

/* name: "W<l>", "b<l>", "GW<l>", "Gb<l>", "H<l>" (layer l output buffer, N x out) */
ORC_API float *orc_gcn_tensor(orc_gcn *g, const char *name, int64_t *count) {
    int l = atoi(name + (name[0] == 'G' ? 2 : 1));
    if (l < 0 || l >= g->L) return NULL;
    orc_layer *y = &g->layers[l];
    if (name[0] == 'W') { *count = y->in * y->out; return y->W; }
    if (name[0] == 'b') { *count = y->out; return y->b; }
    if (name[0] == 'G' && name[1] == 'W') { *count = y->in * y->out; return y->GW; }
    if (name[0] == 'G' && name[1] == 'b') { *count = y->out; return y->Gb; }
    if (name[0] == 'H') { *count = g->N * y->out; return y->AHW; }
    return NULL;
}
