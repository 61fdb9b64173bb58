// dist_row_dn_matrix<r_t> and repl_dn_matrix<r_t> (src/dist_matrix.hpp:394-639).
// Included by dist_matrix.hpp.
#pragma once

namespace mg_gcn {

namespace detail {
// stream `dst` (on dst_ctx's GPU) waits for everything queued so far on stream `src` of src_ctx
inline void stream_wait_stream(const context &dst_ctx, cudaStream_t dst, const context &src_ctx, cudaStream_t src) {
    auto ev = src_ctx.create_event();
    src_ctx.set();
    CHECK_CUDA(cudaEventRecord(ev.get(), src));
    dst_ctx.set();
    CHECK_CUDA(cudaStreamWaitEvent(dst, ev.get(), 0));  // the event may be released once the wait is queued
}
}  // namespace detail

template <typename r_t>
class dist_row_dn_matrix {
    using matrix_t = dn_matrix<r_t>;
    std::vector<matrix_t> As;

   public:
    dist_row_dn_matrix() = default;
    // P blocks of N / P rows each: dense blocks assume N % P == 0 (src/dist_matrix.hpp:427-431)
    dist_row_dn_matrix(const dist_context &ctx, std::size_t N, std::size_t M, std::vector<cuda_ptr<r_t>> buffer = {}) {
        if (N % ctx.size() != 0) throw matrix_error("dist_row_dn_matrix: row count is not a multiple of the GPU count");
        for (std::size_t i = 0; i < ctx.size(); ++i) {
            As.emplace_back(N / ctx.size(), M, i < buffer.size() ? buffer[i] : nullptr);
            pin_managed(As[i].buffer(), As[i].size() * sizeof(r_t), (int)i);  // block i lives on GPU i (no-op for device buffers)
        }
    }
    dist_row_dn_matrix(const std::vector<std::size_t> &p, std::size_t M, std::vector<cuda_ptr<r_t>> buffer = {}) {
        for (std::size_t i = 0; i + 1 < p.size(); ++i) {
            As.emplace_back(p[i + 1] - p[i], M, i < buffer.size() ? buffer[i] : nullptr);
            pin_managed(As[i].buffer(), As[i].size() * sizeof(r_t), (int)i);
        }
    }
    dist_row_dn_matrix(const dist_context &ctx, std::pair<std::size_t, std::size_t> shape)
        : dist_row_dn_matrix(ctx, shape.first, shape.second) {}
    dist_row_dn_matrix(const dist_context &ctx, const matrix_t &A) : dist_row_dn_matrix(ctx, A.n(), A.m()) {
        ctx.sync();
        const r_t *src = A.buffer();
        for (std::size_t i = 0; i < ctx.size(); ++i) {
            std::memcpy(As[i].buffer(), src, As[i].size() * sizeof(r_t));
            prefetch_to_device(As[i].buffer(), As[i].size() * sizeof(r_t), (int)i);
            src += As[i].size();
        }
    }

    void to_dn_matrix(const dist_context &ctx, matrix_t A) const {
        ctx.sync();
        r_t *dst = A.buffer();
        for (std::size_t i = 0; i < As.size(); ++i) {
            std::memcpy(dst, As[i].buffer(), As[i].size() * sizeof(r_t));
            dst += As[i].size();
        }
    }

    // Stage i of the pipelined all-gather: block i lands in bAs[j] on every GPU j, on stream
    // `stream_id` of each GPU (src/dist_matrix.hpp:458-467).  P concurrent copy-engine pulls from
    // the root over NVLink.  Stream semantics are those of the NCCL broadcast it replaces: work
    // queued earlier on the ROOT's stream is visible to every copy, and the root's stream does not
    // proceed until every peer has read the block.
    void bcast(const dist_context &ctx, std::size_t i, dist_row_dn_matrix bAs, int stream_id = 1) {
        const context root = ctx[i];
        const std::size_t bytes = As[i].size() * sizeof(r_t);
        auto ready = root.create_event();
        root.set();
        CHECK_CUDA(cudaEventRecord(ready.get(), root.stream(stream_id)));
        for (std::size_t j = 0; j < size(); ++j) {
            const context cj = ctx[j];
            cj.set();
            if (j != i) CHECK_CUDA(cudaStreamWaitEvent(cj.stream(stream_id), ready.get(), 0));
            CHECK_CUDA(cudaMemcpyPeerAsync(bAs[j].buffer(), (int)j, As[i].buffer(), (int)i, bytes, cj.stream(stream_id)));
            if (j != i) detail::stream_wait_stream(root, root.stream(stream_id), cj, cj.stream(stream_id));
        }
    }

    void init(r_t gain = (r_t)std::sqrt(2 / (1 + 0.01 * 0.01))) {
        for (auto &A : As) A.init(gain / std::sqrt((r_t)size()));
    }
    void zero(const dist_context &ctx) {
        for (std::size_t i = 0; i < As.size(); ++i) As[i].zero(ctx[i]);
    }
    dist_row_dn_matrix copy(const dist_context &ctx, cudaMemcpyKind kind = cudaMemcpyDefault) const {
        dist_row_dn_matrix clone(ctx, n(), m());
        copy_to(ctx, clone, kind);
        return clone;
    }
    void copy_to(const dist_context &ctx, const dist_row_dn_matrix &other, cudaMemcpyKind kind = cudaMemcpyDefault) const {
        for (std::size_t i = 0; i < As.size(); ++i) As[i].copy_to(ctx[i], other[i], kind);
    }

    std::size_t n() const {
        std::size_t N = 0;
        for (auto &A : As) N += A.n();
        return N;
    }
    std::size_t m() const { return As[0].m(); }
    std::pair<std::size_t, std::size_t> shape() const { return {n(), m()}; }
    std::size_t size() const { return As.size(); }
    matrix_t operator[](std::size_t i) const { return As[i]; }
};

// Weights, biases and their gradients: one copy per GPU (src/dist_matrix.hpp:534-639).
template <typename r_t>
class repl_dn_matrix {
    using matrix_t = dn_matrix<r_t>;
    std::vector<matrix_t> As;

    void replicate_from_0(const dist_context &ctx) {
        for (std::size_t j = 1; j < size(); ++j) {
            detail::stream_wait_stream(ctx[j], ctx[j].stream(0), ctx[0], ctx[0].stream(0));
            CHECK_CUDA(cudaMemcpyPeerAsync(As[j].buffer(), (int)j, As[0].buffer(), 0, As[0].size() * sizeof(r_t), ctx[j].stream(0)));
        }
        // write-after-read: GPU 0 may not touch its replica again before every peer has read it (Adam adds
        // weight_decay * W to G_W in place right after the all-reduce of the last layer's gradient; a peer that
        // copied late would see that term and add it a second time)
        for (std::size_t j = 1; j < size(); ++j) detail::stream_wait_stream(ctx[0], ctx[0].stream(0), ctx[j], ctx[j].stream(0));
    }

   public:
    repl_dn_matrix() = default;
    repl_dn_matrix(const dist_context &ctx, std::size_t N, std::size_t M) {
        for (std::size_t i = 0; i < ctx.size(); ++i) {
            As.emplace_back(N, M);
            pin_managed(As[i].buffer(), As[i].size() * sizeof(r_t), (int)i, 'r');  // replica i lives on GPU i
        }
    }
    repl_dn_matrix(const dist_context &ctx, const matrix_t &A) : repl_dn_matrix(ctx, A.n(), A.m()) {
        ctx.sync();
        for (std::size_t j = 0; j < size(); ++j) std::memcpy(As[j].buffer(), A.buffer(), A.size() * sizeof(r_t));
    }

    void to_dn_matrix(const dist_context &ctx, matrix_t &A) const {
        ctx.sync();
        A = As[0];
    }

    // Sum over GPUs, result on every GPU (src/dist_matrix.hpp:587-592).  Added in rank order on
    // GPU 0 and copied out, so every replica holds the same bits.
    void allreduce(const dist_context &ctx) const {
        static_assert(std::is_same<r_t, float>::value, "allreduce is defined for float matrices");
        const context c0 = ctx[0];
        for (std::size_t j = 1; j < size(); ++j) {
            detail::stream_wait_stream(c0, c0.stream(0), ctx[j], ctx[j].stream(0));
            c0.set();
            CHECK_MGGCN(mggcn_axpy_f32(c0.stream(0), As[0].size(), 1.f, As[j].buffer(), As[0].buffer()));
        }
        const_cast<repl_dn_matrix *>(this)->replicate_from_0(ctx);
    }

    std::vector<matrix_t> part2d(std::size_t offset, std::size_t N) const {
        std::vector<matrix_t> out;
        for (auto &A : As) out.emplace_back(N, A.m(), cuda_ptr<r_t>(A.shared_buffer(), A.buffer() + offset * A.m()));
        return out;
    }
    void init(const dist_context &ctx, r_t gain = (r_t)std::sqrt(2 / (1 + 0.01 * 0.01))) {
        if (As.empty()) return;
        ctx.sync();
        As[0].init(gain);
        replicate_from_0(ctx);
    }
    void zero(const dist_context &ctx) {
        for (std::size_t i = 0; i < As.size(); ++i) As[i].zero(ctx[i]);
    }
    repl_dn_matrix copy(const dist_context &ctx, cudaMemcpyKind kind = cudaMemcpyDefault) const {
        repl_dn_matrix clone(ctx, n(), m());
        copy_to(ctx, clone, kind);
        return clone;
    }
    void copy_to(const dist_context &ctx, const repl_dn_matrix &other, cudaMemcpyKind kind = cudaMemcpyDefault) const {
        for (std::size_t i = 0; i < As.size(); ++i) As[i].copy_to(ctx[i], other[i], kind);
    }
    repl_dn_matrix transpose(const dist_context &ctx) const {
        repl_dn_matrix clone(ctx, m(), n());
        for (std::size_t i = 0; i < As.size(); ++i) clone.As[i] = As[i].transpose(ctx[i]);
        return clone;
    }

    std::size_t n() const { return As[0].n(); }
    std::size_t m() const { return As[0].m(); }
    std::pair<std::size_t, std::size_t> shape() const { return {n(), m()}; }
    std::size_t size() const { return As.size(); }
    matrix_t operator[](std::size_t i) const { return As[i]; }
    r_t operator()(std::size_t i) const { return As[0][i]; }
};

}  // namespace mg_gcn
