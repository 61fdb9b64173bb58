// cuda_utils.hpp -- the op layer: the `matmul` overload set (SpMM, staged multi-GPU SpMM, GEMM) and
// the element-wise wrappers around the aggregation.  Surface of the reference's
// src/cuda_utils.hpp:16-172,304-396,470-495; each body is one call into include/mggcn.h.
#pragma once

#include <cassert>

#include "dist_matrix.hpp"

namespace mg_gcn {

// The reference selects a cuSPARSE algorithm here; kept as an ignored tag for source compatibility.
enum spmm_alg_t { SPMM_ALG_DEFAULT = 0 };

namespace detail {
inline std::pair<void *, std::size_t> device_workspace(int device, std::size_t bytes) {
    struct slot {
        cuda_ptr<char> p;
        std::size_t n = 0;
    };
    static std::mutex mu;
    static std::map<int, slot> cache;
    std::lock_guard<std::mutex> g(mu);
    slot &s = cache[device];
    if (bytes > s.n) {
        s.p = cuda_malloc<char>(bytes);  // cudaFree of the old block synchronises: no kernel still reads it
        s.n = bytes;
    }
    return {s.p.get(), s.n};
}
}  // namespace detail

// get_matmul_buffer (src/cuda_utils.hpp:94-147): the new SpMM sizes its own workspace inside the
// matrix's plan, so callers get an empty buffer and may keep passing it around.
template <typename... Args>
cuda_ptr<char> get_matmul_buffer(const context &, Args &&...) { return {}; }
template <typename... Args>
std::vector<cuda_ptr<char>> get_matmul_buffer(const dist_context &ctx, Args &&...) {
    return std::vector<cuda_ptr<char>>(ctx.size());
}

// a1.  C = alpha * A * B + beta * C   (src/cuda_utils.hpp:27-32; cusparseSpMM in the reference)
template <typename x_t, typename v_t, typename r_t>
void matmul(const context &ctx, const csr_matrix<x_t, v_t, r_t> &A, const dn_matrix<r_t> &B, const dn_matrix<r_t> &C,
            const cuda_ptr<char> & /*ext_buffer*/, const r_t alpha, const r_t beta, spmm_alg_t = SPMM_ALG_DEFAULT,
            const mggcn_epilogue *epi = nullptr) {
    assert(A.m() == B.n() && A.n() == C.n() && B.m() == C.m());
    ctx.set();
    auto pw = A.plan(B.m());
    const auto *ci = reinterpret_cast<const std::uint32_t *>(A.indices());
    if (sizeof(x_t) == 8)
        CHECK_MGGCN(mggcn_spmm_csr_f32_i64(ctx.stream(0), A.n(), A.m(), reinterpret_cast<const std::uint64_t *>(A.indptr()), ci,
                                           A.data(), B.buffer(), (std::int64_t)B.m(), C.buffer(), (std::int64_t)C.m(),
                                           (int)B.m(), alpha, beta, epi, pw.first, pw.second));
    else
        CHECK_MGGCN(mggcn_spmm_csr_f32_i32(ctx.stream(0), A.n(), A.m(), reinterpret_cast<const std::uint32_t *>(A.indptr()), ci,
                                           A.data(), B.buffer(), (std::int64_t)B.m(), C.buffer(), (std::int64_t)C.m(),
                                           (int)B.m(), alpha, beta, epi, pw.first, pw.second));
}

// a3.  Staged multi-GPU SpMM with a double-buffered exchange (src/cuda_utils.hpp:57-92):
// stage i moves block i of B into B_bcast[i % 2] on the exchange stream while the compute stream
// multiplies stage i-1; C_j = beta*C_j + sum_i A[j][i] * B_i, accumulated in stage order.
// Timer names follow the reference so its plotting scripts keep working.
template <typename x_t, typename v_t, typename r_t>
void matmul(dist_context ctx, const dist_row_csr_matrix<x_t, v_t, r_t> &A, dist_row_dn_matrix<r_t> B,
            const dist_row_dn_matrix<r_t> &C, const std::vector<cuda_ptr<char>> &ext_buffer,
            std::vector<dist_row_dn_matrix<r_t>> B_bcast, const r_t alpha, const r_t beta, const std::string name = "",
            spmm_alg_t alg = SPMM_ALG_DEFAULT, const mggcn_epilogue *epi = nullptr) {
    const std::size_t P = ctx.size();
    const int comm = ctx.bcast_stream_id();
    const std::size_t nbuf = B_bcast.size();
    static const cuda_ptr<char> none;
    ctx.record(name + "0_matmul-spmm", 0);  // B is complete on every GPU's compute stream
    if (comm != 0) ctx.wait(name + "0_matmul-spmm", comm);
    for (std::size_t i = 0; i < P; ++i) {
        const std::string si = name + std::to_string(i), sn = name + std::to_string(i + 1);
        // WAR: staging buffer i % nbuf was last read by the SpMM of stage i - nbuf
        if (comm != 0 && i >= nbuf) ctx.wait(name + std::to_string(i - nbuf + 1) + "_matmul-spmm", comm);
        ctx.record(si + "_matmul-bcast-start", comm);
        B.bcast(ctx, i, B_bcast[i % nbuf], comm);
        ctx.record(si + "_matmul-bcast-finish", comm);
        if (comm != 0) ctx.wait(si + "_matmul-bcast-finish", 0);  // RAW
        ctx.record(sn + "_matmul-spmm-start", 0);
        for (std::size_t j = 0; j < P; ++j)
            matmul(ctx[j], A[{j, i}], B_bcast[i % nbuf][j], C[j], j < ext_buffer.size() ? ext_buffer[j] : none, alpha,
                   i == 0 ? beta : (r_t)1, alg, i + 1 == P ? epi : nullptr);  // the epilogue sees the complete sum
        ctx.record(sn + "_matmul-spmm", 0);
        ctx.register_timer(si + "_matmul-bcast", si + "_matmul-bcast-start", si + "_matmul-bcast-finish");
        ctx.register_timer(sn + "_matmul-spmm-stage", sn + "_matmul-spmm-start", sn + "_matmul-spmm");
    }
    ctx.register_timer(name + "matmul-spmm", name + "0_matmul-spmm", name + std::to_string(P) + "_matmul-spmm");
}

// a4.  Single staging buffer, everything on the compute stream (src/cuda_utils.hpp:47-55).
template <typename x_t, typename v_t, typename r_t>
void matmul(dist_context ctx, const dist_row_csr_matrix<x_t, v_t, r_t> &A, dist_row_dn_matrix<r_t> B,
            const dist_row_dn_matrix<r_t> &C, const std::vector<cuda_ptr<char>> &ext_buffer,
            dist_row_dn_matrix<r_t> B_bcast, const r_t alpha, const r_t beta, spmm_alg_t alg = SPMM_ALG_DEFAULT) {
    static const cuda_ptr<char> none;
    for (std::size_t i = 0; i < ctx.size(); ++i) {
        B.bcast(ctx, i, B_bcast, 0);
        for (std::size_t j = 0; j < ctx.size(); ++j)
            matmul(ctx[j], A[{j, i}], B_bcast[j], C[j], j < ext_buffer.size() ? ext_buffer[j] : none, alpha,
                   i == 0 ? beta : (r_t)1, alg);
    }
}

// Dense transform, row-major: C = alpha * op(A) * op(B) + beta * C  (src/cuda_utils.hpp:158-172;
// cublasSgemm FP32 in the reference, tcgen05 3xTF32 here).  `epi` optionally fuses bias / activation.
inline void matmul(const context &ctx, const dn_matrix<float> &A, const dn_matrix<float> &B, const dn_matrix<float> &C,
                   const float alpha, const float beta, bool A_T = false, bool B_T = false,
                   const mggcn_epilogue *epi = nullptr) {
    const std::int64_t M = (std::int64_t)(A_T ? A.m() : A.n()), K = (std::int64_t)(A_T ? A.n() : A.m());
    const std::int64_t N = (std::int64_t)(B_T ? B.n() : B.m());
    assert(K == (std::int64_t)(B_T ? B.m() : B.n()) && M == (std::int64_t)C.n() && N == (std::int64_t)C.m());
    ctx.set();
    auto ws = detail::device_workspace((int)ctx.rank, mggcn_gemm_workspace_bytes(A_T, B_T, M, N, K));
    CHECK_MGGCN(mggcn_gemm_f32(ctx.stream(0), A_T, B_T, M, N, K, alpha, A.buffer(), (std::int64_t)A.m(), B.buffer(),
                               (std::int64_t)B.m(), beta, C.buffer(), (std::int64_t)C.m(), epi, ws.first, ws.second));
}

// C (replicated) = alpha * A^T B summed over the row blocks (src/cuda_utils.hpp:304-313)
inline void matmul(const dist_context &ctx, const dist_row_dn_matrix<float> &A, const dist_row_dn_matrix<float> &B,
                   const repl_dn_matrix<float> &C, const float alpha, const float beta) {
    for (std::size_t i = 0; i < ctx.size(); ++i) matmul(ctx[i], A[i], B[i], C[i], alpha, beta, true, false);
    C.allreduce(ctx);
}
// C_i = alpha * A_i * op(B) + beta * C_i with replicated B (src/cuda_utils.hpp:315-324)
inline void matmul(const dist_context &ctx, const dist_row_dn_matrix<float> &A, const repl_dn_matrix<float> &B,
                   const dist_row_dn_matrix<float> &C, const float alpha, const float beta, bool B_T = false) {
    for (std::size_t i = 0; i < ctx.size(); ++i) matmul(ctx[i], A[i], B[i], C[i], alpha, beta, false, B_T);
}

// y += alpha * x   (cublasSaxpy, src/cuda_utils.hpp:326-342; same argument order)
inline void axpy(const context &ctx, const dn_matrix<float> &x, const dn_matrix<float> &y, const float alpha) {
    ctx.set();
    CHECK_MGGCN(mggcn_axpy_f32(ctx.stream(0), x.size(), alpha, x.buffer(), y.buffer()));
}

// y = alpha * x + beta * y   (src/cuda_utils.cu:81-87, wrapper src/cuda_utils.hpp:336-344)
inline void axpby(const context &ctx, const dn_matrix<float> &x, const dn_matrix<float> &y, const float alpha, const float beta) {
    ctx.set();
    CHECK_MGGCN(mggcn_axpby_f32(ctx.stream(0), x.size(), alpha, x.buffer(), beta, y.buffer()));
}

// out = max(in, alpha * in)   (src/cuda_utils.cu:26-31, wrapper src/cuda_utils.hpp:470-483)
inline void leaky_relu_forward(const context &ctx, const dn_matrix<float> &in, const dn_matrix<float> &out, float alpha = 0.01f) {
    ctx.set();
    CHECK_MGGCN(mggcn_leaky_relu_forward_f32(ctx.stream(0), in.buffer(), out.buffer(), in.size(), alpha));
}
// G_out = in > 0 ? G_in : alpha * G_in   (src/cuda_utils.cu:33-38, wrapper :485-495)
inline void leaky_relu_backward(const context &ctx, const dn_matrix<float> &in, const dn_matrix<float> &G_in,
                                const dn_matrix<float> &G_out, float alpha = 0.01f) {
    ctx.set();
    CHECK_MGGCN(mggcn_leaky_relu_backward_f32(ctx.stream(0), in.buffer(), G_in.buffer(), G_out.buffer(), in.size(), alpha));
}
// mat[r, :] = row (discard) or += row   (src/cuda_utils.cu:40-51)
inline void broadcast_rows(const context &ctx, const dn_matrix<float> &row, const dn_matrix<float> &mat, bool discard = true) {
    ctx.set();
    CHECK_MGGCN(mggcn_broadcast_rows_f32(ctx.stream(0), row.buffer(), mat.buffer(), mat.n(), mat.m(), discard ? 1 : 0));
}

}  // namespace mg_gcn
