// gcn.hpp -- the aggregation functors (the drop-in boundary of the hot path), the loss layer and
// the layer and model classes.  Surface of the reference's src/gcn.hpp:13-86 (sparse_linear,
// dist_sparse_linear), :88-296,411-637 (linear, gcn_layer and their dist forms: gcn_layers.hpp),
// :769-823,872-935 (softmax cross-entropy), :937-1056 (gcn, dist_gcn: gcn_model.hpp).
#pragma once

#include "cuda_utils.hpp"

namespace mg_gcn {

// Single-GPU aggregation: forward multiplies by A, backward by A_T (src/gcn.hpp:13-48).
template <typename x_t, typename v_t, typename r_t>
class sparse_linear {
    using csr_t = csr_matrix<x_t, v_t, r_t>;
    using dn_t = dn_matrix<r_t>;
    std::string name;
    csr_t A, A_T;
    cuda_ptr<char> ext_buffer, G_ext_buffer;

   public:
    sparse_linear(const std::string &name, csr_t A, csr_t A_T) : name(name), A(std::move(A)), A_T(std::move(A_T)) {}

    // C = A B (discard) or C += A B; `epi` (optional) is applied to the sum before its only store
    void operator()(context ctx, const dn_t &B, const dn_t &C, bool discard = true, const mggcn_epilogue *epi = nullptr) {
        ctx.record(name + "0_matmul-spmm-start", 0);
        matmul(ctx, A, B, C, ext_buffer, (r_t)1, discard ? (r_t)0 : (r_t)1, SPMM_ALG_DEFAULT, epi);
        ctx.record(name + "0_matmul-spmm-finish", 0);
        ctx.register_timer(name + "0_matmul-spmm", name + "0_matmul-spmm-start", name + "0_matmul-spmm-finish");
    }
    void backward(context ctx, const dn_t &G, const dn_t &G_out, bool discard = true) {
        ctx.record(name + "1_matmul-spmm-start", 0);
        matmul(ctx, A_T, G, G_out, G_ext_buffer, (r_t)1, discard ? (r_t)0 : (r_t)1);
        ctx.record(name + "1_matmul-spmm-finish", 0);
        ctx.register_timer(name + "1_matmul-spmm", name + "1_matmul-spmm-start", name + "1_matmul-spmm-finish");
    }
};

// Multi-GPU aggregation on the row partition (src/gcn.hpp:50-86).  The two staging matrices are
// views over caller-provided buffers (the model shares them between all layers); when none are
// given they are allocated on first use for the width at hand.
template <bool row_partition, typename x_t, typename v_t, typename r_t>
class dist_sparse_linear {
    static_assert(row_partition, "only the row partition is live in the reference (src/main.cpp:171-189)");
    using csr_t = dist_row_csr_matrix<x_t, v_t, r_t>;
    using dn_t = dist_row_dn_matrix<r_t>;
    std::string name;
    csr_t A, A_T;
    std::vector<cuda_ptr<r_t>> bcast_buffer, bcast_buffer2;
    std::vector<cuda_ptr<char>> ext_buffer, G_ext_buffer;
    std::vector<dn_t> B_bcast, G_bcast;

    std::vector<dn_t> staging(const dist_context &ctx, const dn_t &B) {
        std::vector<std::size_t> p{0};
        std::size_t rows = 0;
        for (std::size_t i = 0; i < B.size(); ++i) rows = std::max(rows, B[i].n());
        for (std::size_t i = 0; i < ctx.size(); ++i) p.push_back(p.back() + rows);
        // Staging areas nobody indexes on the host: plain device memory on the GPU that multiplies out of them (the
        // reference's are managed like everything else, src/gcn.hpp:62-70).  MGGCN_STAGING_MANAGED=1 restores that.
        static const bool managed = std::getenv("MGGCN_STAGING_MANAGED") != nullptr;
        std::vector<cuda_ptr<r_t>> buf[2] = {bcast_buffer, bcast_buffer2};  // the caller's, when it passed some
        for (auto &b : buf)
            if (b.empty() && !managed && have_device())
                for (std::size_t i = 0; i < ctx.size(); ++i) {
                    ctx[i].set();
                    b.push_back(cuda_malloc<r_t>(rows * B.m()));  // owned by the dn_matrix views built below
                }
        return {dn_t(p, B.m(), buf[0]), dn_t(p, B.m(), buf[1])};
    }

   public:
    dist_sparse_linear(const std::string &name, csr_t A, csr_t A_T, std::vector<cuda_ptr<r_t>> bcast_buffer = {},
                       std::vector<cuda_ptr<r_t>> bcast_buffer2 = {})
        : name(name), A(std::move(A)), A_T(std::move(A_T)), bcast_buffer(std::move(bcast_buffer)),
          bcast_buffer2(std::move(bcast_buffer2)) {}

    void operator()(const dist_context &ctx, const dn_t &B, const dn_t &C, bool discard = true,
                    const mggcn_epilogue *epi = nullptr) {
        if (B_bcast.empty() || B_bcast[0].m() != B.m()) B_bcast = staging(ctx, B);
        matmul(ctx, A, B, C, ext_buffer, B_bcast, (r_t)1, discard ? (r_t)0 : (r_t)1, name + "0_", SPMM_ALG_DEFAULT, epi);
    }
    void backward(const dist_context &ctx, const dn_t &G, const dn_t &G_out, bool discard = true) {
        if (G_bcast.empty() || G_bcast[0].m() != G.m()) G_bcast = staging(ctx, G);
        matmul(ctx, A_T, G, G_out, G_ext_buffer, G_bcast, (r_t)1, discard ? (r_t)0 : (r_t)1, name + "1_");
    }
};

// Row softmax + cross-entropy + accuracy, and its gradient (softmax - onehot) / N, in ONE kernel
// (the reference: 9 kernels + Sgemm + 2 Sasum, src/gcn.hpp:738-766,769-823).
// operator() blocks (like the reference's ctx.sync() at :928) and returns (mean |log p_y|, accuracy).
template <typename r_t, typename y_t>
class softmax_cross_entropy_loss {
    static_assert(std::is_same<r_t, float>::value && sizeof(y_t) == 4, "float logits, 32-bit labels");
    std::string name;
    bool copy_;
    dn_matrix<r_t> G_;
    cuda_ptr<double> out_;
    cuda_ptr<char> ws_;
    std::size_t ws_bytes_ = 0;

   public:
    explicit softmax_cross_entropy_loss(const std::string &name = "", bool copy = true) : name(name), copy_(copy) {}

    std::pair<r_t, r_t> operator()(context ctx, const dn_matrix<r_t> &H, const dn_matrix<y_t> &Y, std::size_t n_global = 0) {
        ctx.set();
        ctx.record(name + "loss-layer-start", 0);
        G_ = copy_ ? H.copy(ctx) : H;
        const std::size_t need = mggcn_softmax_xent_workspace_bytes((std::int64_t)H.n());
        if (need > ws_bytes_) {
            ws_ = cuda_malloc<char>(need);
            ws_bytes_ = need;
        }
        if (!out_) out_ = cuda_malloc_managed<double>(2);
        const std::size_t N = n_global ? n_global : H.n();
        CHECK_MGGCN(mggcn_softmax_xent_f32(ctx.stream(0), G_.buffer(), (std::int64_t)G_.m(), (std::int64_t)G_.n(), (int)G_.m(),
                                           reinterpret_cast<const std::int32_t *>(Y.buffer()), 1.f / (float)N, out_.get(),
                                           ws_.get(), ws_bytes_));
        ctx.record(name + "loss-layer-finish", 0);
        ctx.register_timer(name + "loss-layer", name + "loss-layer-start", name + "loss-layer-finish");
        ctx.sync();
        return {(r_t)(out_[0] / (double)N), (r_t)(out_[1] / (double)N)};
    }
    dn_matrix<r_t> backward() const { return G_; }
};

}  // namespace mg_gcn

#include "gcn_layers.hpp"
#include "gcn_model.hpp"
