// Host pre-processing of the adjacency: column/row normalisation, transpose and the P x Q block
// split.  Results are bit-exact with the SERIAL forms in the reference
// (src/matrix.hpp:364-388 normalize, :426-453 transpose, src/dist_matrix.hpp:215-259 block split);
// the reference's TBB-parallel forms race, so the serial result is the
// definition.  Everything here is OpenMP-parallel and deterministic for any thread count.
#include <algorithm>
#include <omp.h>
#include <new>
#include <vector>

#include "common.hpp"

using namespace mggcn;

namespace {

// Stable counting sort of the CSR entries by column: returns, for every column, its entries in
// ascending (source row, position) order -- exactly the order the serial transpose visits them.
// Rows are cut into `chunks` contiguous ranges; chunk t owns count[t][c]; the exclusive scan over
// (c major, t minor) gives each chunk its private, ordered output window per column.
struct col_sort {
    int chunks = 1;
    std::vector<int64_t> row_cut;          // chunks + 1
    std::vector<std::vector<uint32_t>> cnt;  // [chunk][cols] -> becomes per-chunk offsets within column
    std::vector<uint64_t> t_rowptr;        // cols + 1
};

void build_col_sort(int64_t rows, int64_t cols, const uint64_t *rowptr, const uint32_t *colidx,
                    col_sort &cs) {
    int T = omp_get_max_threads();
    // memory is chunks * cols * 4 bytes: cap the chunk count for very wide matrices
    int64_t budget = (int64_t(1) << 31) / std::max<int64_t>(1, cols * 4);
    cs.chunks = (int)std::max<int64_t>(1, std::min<int64_t>({(int64_t)T, budget, rows > 0 ? rows : 1}));
    const int C = cs.chunks;
    // balance chunks by nnz
    cs.row_cut.assign(C + 1, 0);
    const uint64_t nnz = rowptr[rows] - rowptr[0];
    for (int t = 1; t < C; ++t) {
        uint64_t target = rowptr[0] + nnz * (uint64_t)t / (uint64_t)C;
        cs.row_cut[t] = std::lower_bound(rowptr, rowptr + rows + 1, target) - rowptr;
        if (cs.row_cut[t] > rows) cs.row_cut[t] = rows;
        if (cs.row_cut[t] < cs.row_cut[t - 1]) cs.row_cut[t] = cs.row_cut[t - 1];
    }
    cs.row_cut[C] = rows;
    cs.cnt.assign(C, {});
#pragma omp parallel for schedule(static, 1) num_threads(C)
    for (int t = 0; t < C; ++t) {
        auto &c = cs.cnt[t];
        c.assign((size_t)cols, 0);
        for (uint64_t e = rowptr[cs.row_cut[t]]; e < rowptr[cs.row_cut[t + 1]]; ++e) c[colidx[e]]++;
    }
    cs.t_rowptr.assign((size_t)cols + 1, 0);
    // per-column totals, then scan (serial scan over cols is cheap next to the nnz passes)
#pragma omp parallel for schedule(static)
    for (int64_t c = 0; c < cols; ++c) {
        uint64_t run = 0;
        for (int t = 0; t < C; ++t) {
            uint32_t k = cs.cnt[t][c];
            cs.cnt[t][c] = (uint32_t)run;  // offset of chunk t inside column c
            run += k;
        }
        cs.t_rowptr[c + 1] = run;
    }
    for (int64_t c = 0; c < cols; ++c) cs.t_rowptr[c + 1] += cs.t_rowptr[c];
}

inline int block_of(const uint32_t *q, int Q, uint32_t c) {
    // the k with q[k] <= c < q[k+1]  (reference: lowerbound_index, src/dist_matrix.hpp:177-185)
    return (int)(std::upper_bound(q, q + Q + 1, c) - q) - 1;
}

}  // namespace

extern "C" {

int mggcn_csr_normalize(int64_t rows, int64_t cols, const uint64_t *rowptr, const uint32_t *colidx,
                        float *vals, int by_column) {
    if (rows < 0 || cols < 0 || !rowptr || (rowptr[rows] && (!colidx || !vals)))
        return fail(MGGCN_E_INVALID, "bad argument");
    if (!by_column) {
#pragma omp parallel for schedule(dynamic, 1024)
        for (int64_t v = 0; v < rows; ++v) {
            float sum = 0;
            for (uint64_t e = rowptr[v]; e < rowptr[v + 1]; ++e) sum += vals[e];
            for (uint64_t e = rowptr[v]; e < rowptr[v + 1]; ++e) vals[e] /= sum;
        }
        return MGGCN_OK;
    }
    // column sums in serial CSR order: gather each column's entries in (row, position) order
    col_sort cs;
    build_col_sort(rows, cols, rowptr, colidx, cs);
    const uint64_t nnz = rowptr[rows];
    std::vector<float> by_col(nnz);
#pragma omp parallel for schedule(static, 1) num_threads(cs.chunks)
    for (int t = 0; t < cs.chunks; ++t) {
        auto &off = cs.cnt[t];
        for (uint64_t e = rowptr[cs.row_cut[t]]; e < rowptr[cs.row_cut[t + 1]]; ++e) {
            uint32_t c = colidx[e];
            by_col[cs.t_rowptr[c] + off[c]++] = vals[e];
        }
    }
    std::vector<float> colsum((size_t)cols);
#pragma omp parallel for schedule(dynamic, 4096)
    for (int64_t c = 0; c < cols; ++c) {
        float s = 0;
        for (uint64_t k = cs.t_rowptr[c]; k < cs.t_rowptr[c + 1]; ++k) s += by_col[k];
        colsum[c] = s;
    }
#pragma omp parallel for schedule(static)
    for (int64_t e = 0; e < (int64_t)nnz; ++e) vals[e] /= colsum[colidx[e]];
    return MGGCN_OK;
}

int mggcn_csr_transpose(int64_t rows, int64_t cols, const uint64_t *rowptr, const uint32_t *colidx,
                        const float *vals, uint64_t *t_rowptr, uint32_t *t_colidx, float *t_vals) {
    if (rows < 0 || cols < 0 || !rowptr || !t_rowptr) return fail(MGGCN_E_INVALID, "bad argument");
    if (rowptr[0] != 0) return fail(MGGCN_E_INVALID, "rowptr[0] must be 0");
    col_sort cs;
    build_col_sort(rows, cols, rowptr, colidx, cs);
    std::copy(cs.t_rowptr.begin(), cs.t_rowptr.end(), t_rowptr);
#pragma omp parallel for schedule(static, 1) num_threads(cs.chunks)
    for (int t = 0; t < cs.chunks; ++t) {
        auto &off = cs.cnt[t];
        for (int64_t r = cs.row_cut[t]; r < cs.row_cut[t + 1]; ++r)
            for (uint64_t e = rowptr[r]; e < rowptr[r + 1]; ++e) {
                uint32_t c = colidx[e];
                uint64_t loc = cs.t_rowptr[c] + off[c]++;
                t_colidx[loc] = (uint32_t)r;
                t_vals[loc] = vals[e];
            }
    }
    return MGGCN_OK;
}

int mggcn_csr_block_counts(int64_t rows, const uint64_t *rowptr, const uint32_t *colidx, int P,
                           const uint32_t *p, int Q, const uint32_t *q, uint64_t **blk_rowptr) {
    if (!rowptr || !p || !q || !blk_rowptr || P <= 0 || Q <= 0) return fail(MGGCN_E_INVALID, "bad argument");
    if ((int64_t)p[P] > rows) return fail(MGGCN_E_INVALID, "row cuts exceed the matrix");
    for (int i = 0; i < P; ++i) {
        const int64_t r0 = p[i], r1 = p[i + 1];
        if (r1 < r0) return fail(MGGCN_E_INVALID, "row cuts must be non-decreasing");
#pragma omp parallel for schedule(dynamic, 512)
        for (int64_t r = r0; r < r1; ++r) {
            uint64_t local[64];
            std::vector<uint64_t> big;
            uint64_t *cnt = local;
            if (Q > 64) { big.assign(Q, 0); cnt = big.data(); } else std::fill(local, local + Q, 0);
            for (uint64_t e = rowptr[r]; e < rowptr[r + 1]; ++e) {
                int j = block_of(q, Q, colidx[e]);
                if (j >= 0 && j < Q) cnt[j]++;
            }
            for (int j = 0; j < Q; ++j) blk_rowptr[(size_t)i * Q + j][r - r0 + 1] = cnt[j];
        }
        for (int j = 0; j < Q; ++j) {
            uint64_t *rp = blk_rowptr[(size_t)i * Q + j];
            rp[0] = 0;
            for (int64_t r = 0; r < r1 - r0; ++r) rp[r + 1] += rp[r];
        }
    }
    return MGGCN_OK;
}

int mggcn_csr_block_fill(int64_t rows, const uint64_t *rowptr, const uint32_t *colidx,
                         const float *vals, int P, const uint32_t *p, int Q, const uint32_t *q,
                         uint64_t *const *blk_rowptr, uint32_t *const *blk_colidx,
                         float *const *blk_vals) {
    if (!rowptr || !p || !q || !blk_rowptr || !blk_colidx || !blk_vals || P <= 0 || Q <= 0)
        return fail(MGGCN_E_INVALID, "bad argument");
    if ((int64_t)p[P] > rows) return fail(MGGCN_E_INVALID, "row cuts exceed the matrix");
    for (int i = 0; i < P; ++i) {
        const int64_t r0 = p[i], r1 = p[i + 1];
#pragma omp parallel for schedule(dynamic, 512)
        for (int64_t r = r0; r < r1; ++r) {
            uint64_t local[64];
            std::vector<uint64_t> big;
            uint64_t *pos = local;
            if (Q > 64) { big.resize(Q); pos = big.data(); }
            for (int j = 0; j < Q; ++j) pos[j] = blk_rowptr[(size_t)i * Q + j][r - r0];
            for (uint64_t e = rowptr[r]; e < rowptr[r + 1]; ++e) {
                const uint32_t c = colidx[e];
                int j = block_of(q, Q, c);
                if (j < 0 || j >= Q) continue;
                const size_t b = (size_t)i * Q + j;
                blk_colidx[b][pos[j]] = c - q[j];
                blk_vals[b][pos[j]] = vals[e];
                pos[j]++;
            }
        }
    }
    return MGGCN_OK;
}

// test/data/prep.py:232-272 (compute_comm): L[i][j] = number of DISTINCT columns of block row i that fall into
// column block j, i.e. how many rows of H_j part i really needs -- against the p[j+1]-p[j] rows the staged
// exchange ships.  One bitmap per block row, set in parallel (idempotent byte stores), counted per column block.
int mggcn_csr_needed_rows(int64_t rows, int64_t cols, const uint64_t *rowptr, const uint32_t *colidx, int P,
                          const uint32_t *p, int Q, const uint32_t *q, uint64_t *needed) {
    if (!rowptr || !p || !q || !needed || P <= 0 || Q <= 0 || rows < 0 || cols < 0) return fail(MGGCN_E_INVALID, "bad argument");
    if ((int64_t)p[P] > rows || (int64_t)q[Q] > cols) return fail(MGGCN_E_INVALID, "cuts exceed the matrix");
    if (rowptr[rows] > 0 && !colidx) return fail(MGGCN_E_INVALID, "colidx is null");
    std::vector<uint8_t> mask;
    try {
        mask.resize((size_t)cols);
    } catch (const std::bad_alloc &) {
        return fail(MGGCN_E_NOMEM, "out of memory");
    }
    for (int i = 0; i < P; ++i) {
        std::fill(mask.begin(), mask.end(), (uint8_t)0);
        uint8_t *m = mask.data();
        const int64_t e0 = (int64_t)rowptr[p[i]], e1 = (int64_t)rowptr[p[i + 1]];
        bool bad = false;
#pragma omp parallel for schedule(static) reduction(|| : bad)
        for (int64_t e = e0; e < e1; ++e) {
            const uint32_t c = colidx[e];
            if ((int64_t)c < cols) m[c] = 1;
            else bad = true;
        }
        if (bad) return fail(MGGCN_E_INVALID, "column index outside the matrix");
        for (int j = 0; j < Q; ++j) {
            uint64_t cnt = 0;
            const int64_t c0 = q[j], c1 = q[j + 1];
#pragma omp parallel for schedule(static) reduction(+ : cnt)
            for (int64_t c = c0; c < c1; ++c) cnt += m[c];
            needed[(size_t)i * Q + j] = cnt;
        }
    }
    return MGGCN_OK;
}

}  // extern "C"
