// gcn<x_t,v_t,r_t> and dist_gcn<true,...>: the model classes of src/gcn.hpp:937-1056 over the epoch
// engine of the C-ABI (mggcn_engine_*).  Included by gcn.hpp.
//
// Same call sequence as the reference's epoch loop (src/main.cpp:113-133,153-169):
//     auto [loss, acc] = G.train_forward(ctx, X, Y);  G.backward(ctx);  G.adam_update(ctx, ...);
// plus one extension, train_epoch(), which runs the three steps as one replayed CUDA graph.
// Not provided: residual layers (off in the reference CLI), the SGD `update` (unused by the CLI),
// inference-only operator() (only the reference's disabled test calls it).
#pragma once

namespace mg_gcn {

namespace detail {

// owning handle of one engine (= one GPU's share of the model)
struct engine_box {
    mggcn_engine_t *e = nullptr;
    const void *x_seen = nullptr, *y_seen = nullptr;
    ~engine_box() {
        if (e) mggcn_engine_destroy(e);
    }
};

template <typename csr_t>
void upload_block_row(mggcn_engine_t *e, int which, const std::vector<csr_t> &blocks) {
    const std::size_t P = blocks.size();
    std::vector<std::vector<std::uint64_t>> rp(P);
    std::vector<const std::uint64_t *> rpp(P);
    std::vector<const std::uint32_t *> cip(P);
    std::vector<const float *> vap(P);
    for (std::size_t j = 0; j < P; ++j) {
        const csr_t &b = blocks[j];
        rp[j].resize((std::size_t)b.n() + 1);
        for (std::size_t r = 0; r <= b.n(); ++r) rp[j][r] = (std::uint64_t)b.indptr()[r];
        rpp[j] = rp[j].data();
        cip[j] = reinterpret_cast<const std::uint32_t *>(b.indices());
        vap[j] = b.data();
    }
    CHECK_MGGCN(mggcn_engine_set_adjacency(e, which, rpp.data(), cip.data(), vap.data()));
}

inline dn_matrix<float> fetch(mggcn_engine_t *e, const std::string &name, std::size_t rows, std::size_t cols) {
    dn_matrix<float> out(rows, cols);
    std::size_t count = 0;
    CHECK_MGGCN(mggcn_engine_get(e, name.c_str(), out.buffer(), out.size(), &count));
    return out;
}

}  // namespace detail

// What `G.layers()[l]` exposes in the reference: W(), b(), GW(), Gb() (src/gcn.hpp:503-517,626-637).
class gcn_layer_view {
    mggcn_engine_t *e_;
    int l_;
    std::size_t in_, out_;

   public:
    gcn_layer_view(mggcn_engine_t *e, int l, std::size_t in, std::size_t out) : e_(e), l_(l), in_(in), out_(out) {}
    dn_matrix<float> W() const { return detail::fetch(e_, "W" + std::to_string(l_), in_, out_); }
    dn_matrix<float> b() const { return detail::fetch(e_, "b" + std::to_string(l_), 1, out_); }
    dn_matrix<float> GW() const { return detail::fetch(e_, "GW" + std::to_string(l_), in_, out_); }
    dn_matrix<float> Gb() const { return detail::fetch(e_, "Gb" + std::to_string(l_), 1, out_); }
};

struct gcn_options {
    bool exact_grads = false;  // false = the reference's gradient quirks Q1/Q2
    bool use_graph = true;     // train_epoch() replays a captured CUDA graph
    bool timers = false;       // eager execution with the reference's named timers
    float lrelu_slope = 0.01f;
};

template <typename x_t, typename v_t, typename r_t>
class gcn {
    static_assert(std::is_same<r_t, float>::value, "the model is fp32, like the reference CLI");
    using csr_t = csr_matrix<x_t, v_t, r_t>;
    csr_t A_, A_T_;
    std::vector<std::size_t> sizes_;
    gcn_options opt_;
    std::shared_ptr<detail::engine_box> box_ = std::make_shared<detail::engine_box>();

    mggcn_engine_t *engine(const context &ctx) {
        if (box_->e) return box_->e;
        std::vector<std::int64_t> sz(sizes_.begin(), sizes_.end());
        const std::uint32_t cuts[2] = {0, (std::uint32_t)A_.n()};
        mggcn_engine_config cfg{};
        cfg.world = 1;
        cfg.device = (int)ctx.rank;
        cfg.num_layers = (int)sz.size() - 1;
        cfg.sizes = sz.data();
        cfg.n_global = A_.n();
        cfg.p = cuts;
        cfg.overlap = 1;
        cfg.exact_grads = opt_.exact_grads;
        cfg.use_graph = opt_.use_graph && !opt_.timers;
        cfg.timers = opt_.timers;
        cfg.lrelu_slope = opt_.lrelu_slope;
        CHECK_MGGCN(mggcn_engine_create(&cfg, &box_->e));
        detail::upload_block_row<csr_t>(box_->e, 0, {A_T_});  // forward operand (src/gcn.hpp:947-954)
        detail::upload_block_row<csr_t>(box_->e, 1, {A_});    // backward operand
        return box_->e;
    }
    void inputs(mggcn_engine_t *e, const dn_matrix<r_t> &H, const dn_matrix<std::int32_t> &Y) {
        const bool nx = H.buffer() != box_->x_seen, ny = Y.buffer() != box_->y_seen;
        if (nx || ny) CHECK_MGGCN(mggcn_engine_set_inputs(e, nx ? H.buffer() : nullptr, ny ? Y.buffer() : nullptr, 0));
        box_->x_seen = H.buffer();
        box_->y_seen = Y.buffer();
    }
    std::pair<r_t, r_t> loss(mggcn_engine_t *e) {
        double s = 0, c = 0;
        CHECK_MGGCN(mggcn_engine_loss(e, &s, &c));
        return {(r_t)(s / (double)A_.n()), (r_t)(c / (double)A_.n())};
    }

   public:
    // normalises by column and transposes INSIDE, like the reference (src/gcn.hpp:946-955)
    gcn(csr_t A, std::vector<std::size_t> sizes, bool residual_layer = false, gcn_options opt = {})
        : A_(std::move(A)), sizes_(std::move(sizes)), opt_(opt) {
        if (residual_layer) throw std::invalid_argument("residual layers are not built (off in the reference CLI)");
        A_.normalize(true);
        A_T_ = A_.transpose();
    }

    std::pair<r_t, r_t> train_forward(const context &ctx, const dn_matrix<r_t> &H, const dn_matrix<std::int32_t> &Y) {
        auto *e = engine(ctx);
        inputs(e, H, Y);
        CHECK_MGGCN(mggcn_engine_train_forward(e));
        return loss(e);
    }
    void backward(const context &ctx) { CHECK_MGGCN(mggcn_engine_backward(engine(ctx))); }
    void adam_update(const context &ctx, r_t lr, r_t beta1, r_t beta2, r_t wd, r_t eps) {
        CHECK_MGGCN(mggcn_engine_adam_update(engine(ctx), lr, beta1, beta2, wd, eps));
    }
    // forward + loss + backward + Adam as one graph replay; returns this epoch's (loss, accuracy)
    std::pair<r_t, r_t> train_epoch(const context &ctx, const dn_matrix<r_t> &H, const dn_matrix<std::int32_t> &Y,
                                    r_t lr, r_t beta1, r_t beta2, r_t wd, r_t eps) {
        auto *e = engine(ctx);
        inputs(e, H, Y);
        CHECK_MGGCN(mggcn_engine_epoch(e, lr, beta1, beta2, wd, eps));
        return loss(e);
    }
    void update(const context &, r_t, r_t) { throw std::logic_error("SGD update is not built; use adam_update"); }
    // waits for everything the model has queued (the model runs on its own streams, not on ctx's)
    void sync() {
        if (box_->e) CHECK_MGGCN(mggcn_engine_sync(box_->e));
    }

    std::vector<gcn_layer_view> layers() const {
        std::vector<gcn_layer_view> out;
        for (std::size_t l = 0; l + 1 < sizes_.size(); ++l) out.emplace_back(box_->e, (int)l, sizes_[l], sizes_[l + 1]);
        return out;
    }
    void dump_timers(std::ostream &out, const std::string &prefix) {
        if (!box_->e) return;
        std::size_t n = 0;
        CHECK_MGGCN(mggcn_engine_dump_timers(box_->e, prefix.c_str(), nullptr, 0, &n));
        std::string buf(n + 1, '\0');
        CHECK_MGGCN(mggcn_engine_dump_timers(box_->e, prefix.c_str(), buf.data(), buf.size(), &n));
        out.write(buf.data(), (std::streamsize)n);
    }
    mggcn_engine_t *handle(const context &ctx) { return engine(ctx); }
};

}  // namespace mg_gcn

#include "gcn_dist_model.hpp"
