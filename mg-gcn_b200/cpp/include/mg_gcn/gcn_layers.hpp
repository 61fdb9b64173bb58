// gcn_layers.hpp -- the layer-level classes around the aggregation: linear, dist_row_linear, gcn_layer,
// dist_gcn_layer.  Surface of the reference's src/gcn.hpp:88-296 (dense transform + its SGD / Adam
// steps) and :411-637 (one graph-convolution layer = aggregation + transform + LeakyReLU, in the cheaper
// of the two orders).  Included by gcn.hpp.
//
// The model classes (gcn_model.hpp) run a whole epoch inside the engine; these classes are for callers
// that compose layers themselves, as the reference's model constructors do (src/gcn.hpp:946-957).  Each
// body is a handful of calls into the op layer (cuda_utils.hpp), with three fusions the reference does
// not have: the bias pre-fill is the GEMM's epilogue instead of broadcast_rows + beta = 1, LeakyReLU is
// the epilogue of whichever product comes last in the layer, and Adam is one kernel per tensor.
//
// Gradient quirks of the reference are kept by default: a layer built with
// backward_spmm = false skips the backward aggregation (Q1), and an aggregation-first layer takes its
// weight gradient against the layer INPUT rather than the aggregated input (Q2, `lin.setX(H)`).
#pragma once

#include <optional>

namespace mg_gcn {

// ---- element-wise wrappers over every GPU of a dist_context (src/cuda_utils.hpp:336-396,497-520) ----
template <typename D>
void axpy(const dist_context &ctx, const D &x, const D &y, const float alpha) {
    for (std::size_t i = 0; i < ctx.size(); ++i) axpy(ctx[i], x[i], y[i], alpha);
}
template <typename D>
void axpby(const dist_context &ctx, const D &x, const D &y, const float alpha, const float beta) {
    for (std::size_t i = 0; i < ctx.size(); ++i) axpby(ctx[i], x[i], y[i], alpha, beta);
}
inline void leaky_relu_forward(const dist_context &ctx, const dist_row_dn_matrix<float> &in,
                               const dist_row_dn_matrix<float> &out, float alpha = 0.01f) {
    for (std::size_t i = 0; i < ctx.size(); ++i) leaky_relu_forward(ctx[i], in[i], out[i], alpha);
}
inline void leaky_relu_backward(const dist_context &ctx, const dist_row_dn_matrix<float> &in,
                                const dist_row_dn_matrix<float> &G_in, const dist_row_dn_matrix<float> &G_out,
                                float alpha = 0.01f) {
    for (std::size_t i = 0; i < ctx.size(); ++i) leaky_relu_backward(ctx[i], in[i], G_in[i], G_out[i], alpha);
}
inline void broadcast_rows(const dist_context &ctx, const repl_dn_matrix<float> &row, const dist_row_dn_matrix<float> &mat,
                           bool discard = true) {
    for (std::size_t i = 0; i < ctx.size(); ++i) broadcast_rows(ctx[i], row[i], mat[i], discard);
}

namespace detail {

// G_b = 1^T G  (the reference multiplies by a row of ones, src/gcn.hpp:126-131)
inline void column_sums(const context &ctx, const dn_matrix<float> &G, const dn_matrix<float> &G_b) {
    ctx.set();
    auto ws = device_workspace((int)ctx.rank, mggcn_colsum_workspace_bytes((std::int64_t)G.n(), (std::int64_t)G.m()));
    CHECK_MGGCN(mggcn_colsum_f32(ctx.stream(0), (std::int64_t)G.n(), (std::int64_t)G.m(), G.buffer(), (std::int64_t)G.m(),
                                 G_b.buffer(), ws.first, ws.second));
}

// XW = [XW +] b + X W, optionally through LeakyReLU: one GEMM (src/gcn.hpp:116-123 is broadcast_rows + Sgemm)
inline void affine(const context &ctx, const dn_matrix<float> &X, const dn_matrix<float> &W, const dn_matrix<float> &b,
                   const dn_matrix<float> &XW, bool discard, bool activation, float slope) {
    mggcn_epilogue epi{};
    epi.kind = activation ? MGGCN_EPI_LEAKY_RELU : MGGCN_EPI_NONE;
    epi.slope = slope;
    epi.bias = b.buffer();
    matmul(ctx, X, W, XW, 1.f, discard ? 0.f : 1.f, false, false, &epi);
}

// One Adam step on a tensor; `decay` = weights (the reference decays W only, src/gcn.hpp:163,270)
inline void adam_step(const context &ctx, const dn_matrix<float> &p, const dn_matrix<float> &g, const dn_matrix<float> &m,
                      const dn_matrix<float> &v, float lr, float beta1, float beta2, float wd, float eps, float bc1,
                      float bc2, bool decay) {
    ctx.set();
    CHECK_MGGCN(mggcn_adam_f32(ctx.stream(0), p.size(), p.buffer(), g.buffer(), m.buffer(), v.buffer(), lr, beta1, beta2, wd,
                               eps, bc1, bc2, decay ? p.size() : 0));
}

}  // namespace detail

// -------------------------------------------------------------------------------------------------
// linear: XW = X W + b, its gradients and its optimiser steps (src/gcn.hpp:88-189).
// -------------------------------------------------------------------------------------------------
template <typename r_t>
class linear {
    static_assert(std::is_same<r_t, float>::value, "fp32, like the reference CLI");
    using dn_t = dn_matrix<r_t>;
    std::string name;
    dn_t W, G_W, mW, vW;
    dn_t b, G_b, mb, vb;
    dn_t X;
    bool backward_out;
    std::size_t step = 0;

   public:
    linear(std::string name, std::size_t in, std::size_t out, bool backward_out = true)
        : name(std::move(name)), W(in, out), G_W(in, out), b(1, out), G_b(1, out), backward_out(backward_out) {
        W.init();
        b.init(std::sqrt((r_t)1.0 / 3));
    }

    void setX(dn_t new_X) { X = new_X; }

    // XW = b + X W (discard) or XW += b + X W; `activation` folds LeakyReLU into the same pass
    void operator()(context ctx, dn_t X, dn_t XW, bool discard = true, bool activation = false, r_t slope = (r_t)0.01) {
        ctx.record(name + "0_0_matmul-gemm", 0);
        detail::affine(ctx, X, W, b, XW, discard, activation, slope);
        ctx.record(name + "0_1_matmul-gemm", 0);
        ctx.register_timer(name + "0_matmul-gemm", name + "0_0_matmul-gemm", name + "0_1_matmul-gemm");
        this->X = X;
    }

    // G_b = 1^T G;  G_W = X^T G;  G_out = G W^T (discard) or += (src/gcn.hpp:125-139)
    void backward(context ctx, dn_t G, dn_t G_out, bool discard = true) {
        ctx.record(name + "1_0_matmul-gemm", 0);
        detail::column_sums(ctx, G, G_b);
        ctx.record(name + "1_1_matmul-gemm", 0);
        matmul(ctx, X, G, G_W, (r_t)1, (r_t)0, true);
        ctx.record(name + "1_2_matmul-gemm", 0);
        if (backward_out) matmul(ctx, G, W, G_out, (r_t)1, discard ? (r_t)0 : (r_t)1, false, true);
        ctx.record(name + "1_3_matmul-gemm", 0);
        ctx.register_timer(name + "1_matmul-gemm", name + "1_0_matmul-gemm", name + "1_3_matmul-gemm");
    }

    // SGD: W = (1 - wd) W - lr G_W;  b -= lr G_b  (src/gcn.hpp:141-144)
    void update(const context ctx, const r_t lr, const r_t weight_decay) {
        axpby(ctx, G_W, W, -lr, 1 - weight_decay);
        axpy(ctx, G_b, b, -lr);
    }

    void adam_update(context ctx, const r_t lr, const r_t beta1, const r_t beta2, const r_t weight_decay, const r_t eps) {
        if (mW.shape() != W.shape()) {
            mW = dn_t(W.shape());
            vW = dn_t(W.shape());
            mb = dn_t(b.shape());
            vb = dn_t(b.shape());
            for (dn_t *t : {&mW, &vW, &mb, &vb}) t->zero(ctx);
            step = 0;
        }
        step += 1;
        const r_t bc1 = 1 - std::pow(beta1, step), bc2 = 1 - std::pow(beta2, step);
        ctx.record(name + "0_adam-update", 0);
        detail::adam_step(ctx, W, G_W, mW, vW, lr, beta1, beta2, weight_decay, eps, bc1, bc2, true);
        detail::adam_step(ctx, b, G_b, mb, vb, lr, beta1, beta2, weight_decay, eps, bc1, bc2, false);
        ctx.record(name + "1_adam-update", 0);
        ctx.register_timer(name + "adam-update", name + "0_adam-update", name + "1_adam-update");
    }

    auto get_b() { return b; }
    auto get_W() { return W; }
    auto get_G_W() { return G_W; }
    auto get_G_b() { return G_b; }
};

// -------------------------------------------------------------------------------------------------
// dist_row_linear: the same on row blocks with replicated parameters; gradients are summed over the
// GPUs (src/gcn.hpp:191-296).
// -------------------------------------------------------------------------------------------------
template <typename r_t>
class dist_row_linear {
    static_assert(std::is_same<r_t, float>::value, "fp32, like the reference CLI");
    using dn_t = dist_row_dn_matrix<r_t>;
    using rdn_t = repl_dn_matrix<r_t>;
    std::string name;
    rdn_t W, G_W, mW, vW;
    rdn_t b, G_b, mb, vb;
    dn_t X;
    bool backward_out;
    std::size_t step = 0;

   public:
    dist_row_linear(const dist_context ctx, std::string name, std::size_t in, std::size_t out, bool backward_out = true)
        : name(std::move(name)), W(ctx, in, out), G_W(ctx, in, out), b(ctx, 1, out), G_b(ctx, 1, out), backward_out(backward_out) {
        W.init(ctx);
        b.init(ctx, std::sqrt((r_t)1.0 / 3));
    }

    void setX(dn_t new_X) { X = new_X; }

    void operator()(dist_context ctx, dn_t X, dn_t XW, bool discard = true, bool activation = false, r_t slope = (r_t)0.01) {
        ctx.record(name + "0_0_matmul-gemm", 0);
        for (std::size_t i = 0; i < ctx.size(); ++i) detail::affine(ctx[i], X[i], W[i], b[i], XW[i], discard, activation, slope);
        ctx.record(name + "0_1_matmul-gemm", 0);
        ctx.register_timer(name + "0_matmul-gemm", name + "0_0_matmul-gemm", name + "0_1_matmul-gemm");
        this->X = X;
    }

    void backward(dist_context ctx, dn_t G, dn_t G_out, bool discard = true) {
        ctx.record(name + "1_0_matmul-gemm", 0);
        for (std::size_t i = 0; i < ctx.size(); ++i) detail::column_sums(ctx[i], G[i], G_b[i]);
        G_b.allreduce(ctx);
        ctx.record(name + "1_1_matmul-gemm", 0);
        matmul(ctx, X, G, G_W, (r_t)1, (r_t)0);  // X^T G on every block, then the all-reduce
        ctx.record(name + "1_2_matmul-gemm", 0);
        if (backward_out) matmul(ctx, G, W, G_out, (r_t)1, discard ? (r_t)0 : (r_t)1, true);
        ctx.record(name + "1_3_matmul-gemm", 0);
        ctx.register_timer(name + "1_matmul-gemm", name + "1_0_matmul-gemm", name + "1_3_matmul-gemm");
    }

    void update(const dist_context ctx, const r_t lr, const r_t weight_decay) {
        axpby(ctx, G_W, W, -lr, 1 - weight_decay);
        axpy(ctx, G_b, b, -lr);
    }

    // every replica takes the same step from the same (all-reduced) gradient, so they stay bit-identical
    void adam_update(dist_context ctx, const r_t lr, const r_t beta1, const r_t beta2, const r_t weight_decay, const r_t eps) {
        if (mW.size() != W.size()) {
            mW = rdn_t(ctx, W.n(), W.m());
            vW = rdn_t(ctx, W.n(), W.m());
            mb = rdn_t(ctx, b.n(), b.m());
            vb = rdn_t(ctx, b.n(), b.m());
            for (rdn_t *t : {&mW, &vW, &mb, &vb}) t->zero(ctx);
            step = 0;
        }
        step += 1;
        const r_t bc1 = 1 - std::pow(beta1, step), bc2 = 1 - std::pow(beta2, step);
        ctx.record(name + "0_adam-update", 0);
        for (std::size_t i = 0; i < ctx.size(); ++i) {
            detail::adam_step(ctx[i], W[i], G_W[i], mW[i], vW[i], lr, beta1, beta2, weight_decay, eps, bc1, bc2, true);
            detail::adam_step(ctx[i], b[i], G_b[i], mb[i], vb[i], lr, beta1, beta2, weight_decay, eps, bc1, bc2, false);
        }
        ctx.record(name + "1_adam-update", 0);
        ctx.register_timer(name + "adam-update", name + "0_adam-update", name + "1_adam-update");
    }

    auto get_b() { return b; }
    auto get_W() { return W; }
    auto get_G_W() { return G_W; }
    auto get_G_b() { return G_b; }
};

namespace detail {

// The body shared by gcn_layer and dist_gcn_layer (src/gcn.hpp:434-500 and :556-623 are the same text
// over different container types).  Traits supply the containers; everything else is overload resolution.
template <typename ctx_t, typename agg_t, typename lin_t, typename dn_t>
class gcn_layer_body {
   protected:
    std::string name;
    agg_t A;
    lin_t lin;
    std::optional<lin_t> res_lin;
    dn_t HW, AHW, G_HW, G_out;  // HW / G_HW share HW_buffer, AHW / G_out share AHW_buffer (src/gcn.hpp:419-424)
    bool activation, residual_layer, backward_spmm;
    dn_t H;
    float slope = 0.01f;

    gcn_layer_body(std::string name, agg_t A, lin_t lin, std::optional<lin_t> res_lin, dn_t HW, dn_t AHW, dn_t G_HW, dn_t G_out,
                   bool activation, bool residual_layer, bool backward_spmm)
        : name(std::move(name)), A(std::move(A)), lin(std::move(lin)), res_lin(std::move(res_lin)), HW(std::move(HW)),
          AHW(std::move(AHW)), G_HW(std::move(G_HW)), G_out(std::move(G_out)), activation(activation),
          residual_layer(residual_layer), backward_spmm(backward_spmm) {}

   public:
    // H is m x in.  The narrower side goes through the aggregation (src/gcn.hpp:437-444).
    auto operator()(ctx_t ctx, dn_t H) {
        this->H = H;
        mggcn_epilogue act{};
        act.kind = MGGCN_EPI_LEAKY_RELU;
        act.slope = slope;
        if (HW.m() == AHW.m()) {
            lin(ctx, H, HW);
            A(ctx, HW, AHW, true, activation ? &act : nullptr);
        } else {
            A(ctx, H, HW);
            lin(ctx, HW, AHW, true, activation, slope);
        }
        if (res_lin)
            (*res_lin)(ctx, H, AHW, false);
        else if (residual_layer)
            axpy(ctx, H, AHW, 1.f);
        return AHW;
    }

    auto backward(ctx_t ctx, dn_t G) {
        auto T = G;
        if (activation) {
            ctx.record(name + "1_0_activation", 0);
            leaky_relu_backward(ctx, AHW, G, AHW, slope);
            ctx.record(name + "1_1_activation", 0);
            ctx.register_timer(name + "1_activation", name + "1_0_activation", name + "1_1_activation");
            T = AHW;
        }
        if (HW.m() == AHW.m()) {
            if (backward_spmm)
                A.backward(ctx, T, G_HW);
            else
                G_HW = T;  // Q1
            lin.backward(ctx, G_HW, G_out);
        } else {
            lin.setX(H);  // Q2
            lin.backward(ctx, T, G_HW);
            if (backward_spmm)
                A.backward(ctx, G_HW, G_out);
            else
                G_out = G_HW;
        }
        if (res_lin)
            res_lin->backward(ctx, G, G_out, false);
        else if (residual_layer)
            axpy(ctx, G, G_out, 1.f);
        return G_out;
    }

    void update(const ctx_t ctx, const float lr, const float weight_decay) {
        lin.update(ctx, lr, weight_decay);
        if (res_lin) res_lin->update(ctx, lr, weight_decay);
    }
    void adam_update(const ctx_t ctx, const float lr, const float beta1, const float beta2, const float weight_decay,
                     const float eps) {
        lin.adam_update(ctx, lr, beta1, beta2, weight_decay, eps);
        if (res_lin) res_lin->adam_update(ctx, lr, beta1, beta2, weight_decay, eps);
    }

    auto b() { return lin.get_b(); }
    auto W() { return lin.get_W(); }
    auto GW() { return lin.get_G_W(); }
    auto Gb() { return lin.get_G_b(); }
};

inline std::vector<cuda_ptr<float>> per_gpu_buffers(const dist_context &ctx, std::size_t elems_per_gpu) {
    std::vector<cuda_ptr<float>> t;
    for (std::size_t i = 0; i < ctx.size(); ++i) {
        ctx[i].set();
        t.emplace_back(cuda_malloc<float>(elems_per_gpu));
    }
    return t;
}

}  // namespace detail

// -------------------------------------------------------------------------------------------------
// gcn_layer (src/gcn.hpp:411-518)
// -------------------------------------------------------------------------------------------------
template <typename x_t, typename v_t, typename r_t>
class gcn_layer : public detail::gcn_layer_body<context, sparse_linear<x_t, v_t, r_t>, linear<r_t>, dn_matrix<r_t>> {
    using base = detail::gcn_layer_body<context, sparse_linear<x_t, v_t, r_t>, linear<r_t>, dn_matrix<r_t>>;
    using csr_t = csr_matrix<x_t, v_t, r_t>;
    using dn_t = dn_matrix<r_t>;

    gcn_layer(std::string name, csr_t A, csr_t A_T, std::size_t in, std::size_t out, bool activation, bool residual_layer,
              bool backward_spmm, cuda_ptr<r_t> HW_buffer, cuda_ptr<r_t> AHW_buffer)
        : base(name, sparse_linear<x_t, v_t, r_t>(name, A, A_T), linear<r_t>(name, in, out, backward_spmm),
               in == out || !residual_layer ? std::nullopt : std::make_optional(linear<r_t>(name, in, out, backward_spmm)),
               dn_t(A.m(), std::min(in, out), HW_buffer), dn_t(A.n(), out, AHW_buffer),
               dn_t(A_T.n(), std::min(in, out), HW_buffer), dn_t(A_T.n(), in, AHW_buffer), activation, residual_layer,
               backward_spmm) {}

   public:
    // HW_buffer may be shared between layers (the models pass one buffer of the widest layer, src/gcn.hpp:942-952)
    gcn_layer(std::string name, csr_t A, csr_t A_T, std::size_t in, std::size_t out, bool activation, bool residual_layer = false,
              bool backward_spmm = true, cuda_ptr<r_t> HW_buffer = nullptr)
        : gcn_layer(name, A, A_T, in, out, activation, residual_layer, backward_spmm,
                    HW_buffer ? HW_buffer
                              : cuda_malloc_managed<r_t>(std::max<std::size_t>(A.m(), A_T.n()) * std::min(in, out)),
                    cuda_malloc_managed<r_t>(std::max((std::size_t)A.n() * out, (std::size_t)A_T.n() * in))) {}
};

// -------------------------------------------------------------------------------------------------
// dist_gcn_layer on the row partition (src/gcn.hpp:520-637)
// -------------------------------------------------------------------------------------------------
template <bool row_partition, typename x_t, typename v_t, typename r_t>
class dist_gcn_layer : public detail::gcn_layer_body<dist_context, dist_sparse_linear<row_partition, x_t, v_t, r_t>,
                                                     dist_row_linear<r_t>, dist_row_dn_matrix<r_t>> {
    static_assert(row_partition, "only the row partition is live in the reference (src/main.cpp:171-189)");
    using base = detail::gcn_layer_body<dist_context, dist_sparse_linear<row_partition, x_t, v_t, r_t>, dist_row_linear<r_t>,
                                        dist_row_dn_matrix<r_t>>;
    using csr_t = dist_row_csr_matrix<x_t, v_t, r_t>;
    using dn_t = dist_row_dn_matrix<r_t>;
    using bufs_t = std::vector<cuda_ptr<r_t>>;

    dist_gcn_layer(const dist_context &ctx, std::string name, csr_t A, csr_t A_T, std::size_t in, std::size_t out,
                   bool activation, bool residual_layer, bool backward_spmm, bufs_t HW_buffer, bufs_t AHW_buffer,
                   bufs_t bcast_buffer, bufs_t bcast_buffer2)
        : base(name, dist_sparse_linear<row_partition, x_t, v_t, r_t>(name, A, A_T, bcast_buffer, bcast_buffer2),
               dist_row_linear<r_t>(ctx, name, in, out, backward_spmm),
               in == out || !residual_layer ? std::nullopt
                                            : std::make_optional(dist_row_linear<r_t>(ctx, name, in, out, backward_spmm)),
               dn_t(ctx, A.m(), std::min(in, out), HW_buffer), dn_t(ctx, A.n(), out, AHW_buffer),
               dn_t(ctx, A_T.n(), std::min(in, out), HW_buffer), dn_t(ctx, A_T.n(), in, AHW_buffer), activation,
               residual_layer, backward_spmm) {}

   public:
    dist_gcn_layer(const dist_context ctx, std::string name, csr_t A, csr_t A_T, std::size_t in, std::size_t out,
                   bool activation, bool residual_layer = false, bool backward_spmm = true, bufs_t HW_buffer = {},
                   bufs_t bcast_buffer = {}, bufs_t bcast_buffer2 = {})
        : dist_gcn_layer(ctx, name, A, A_T, in, out, activation, residual_layer, backward_spmm,
                         !HW_buffer.empty() ? HW_buffer
                                            : detail::per_gpu_buffers(ctx, std::max<std::size_t>(A.m(), A_T.n()) *
                                                                               std::min(in, out) / ctx.size()),
                         detail::per_gpu_buffers(ctx, std::max((std::size_t)A.n() * out, (std::size_t)A_T.n() * in) / ctx.size()),
                         bcast_buffer, bcast_buffer2) {}
};

}  // namespace mg_gcn
