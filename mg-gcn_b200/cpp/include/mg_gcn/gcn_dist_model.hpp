// dist_gcn<true, x_t, v_t, r_t>: the multi-GPU model of src/gcn.hpp:997-1056 -- one engine per GPU,
// all driven from this process (the reference is single-process too).  Included by gcn_model.hpp.
//
// Exchange back-end: peer memory, grouped form (MGGCN_COMM_P2P_GROUPED: plain SpMM launches over groups of
// stages, ordered behind the transfers with stream events).  The single-launch gathered form (MGGCN_COMM_P2P) is for one process per GPU: with
// every engine in ONE process, peer copies between two devices of the same process queue behind kernels that
// are still running on the other device, and a kernel that waits on the device for such a copy can stall until
// the barrier timeout.  The engines are wired to each other in-process
// (mggcn_engine_comm_handle / _connect), every GPU pulls the feature block it needs at stage s
// straight out of its owner's arena over NVLink, and weight gradients are summed in rank order.
#pragma once

namespace mg_gcn {

template <bool row_partition, typename x_t, typename v_t, typename r_t>
class dist_gcn {
    static_assert(row_partition, "only the row partition is live in the reference (src/main.cpp:171-189)");
    static_assert(std::is_same<r_t, float>::value, "the model is fp32, like the reference CLI");
    using csr_t = dist_row_csr_matrix<x_t, v_t, r_t>;
    using block_t = csr_matrix<x_t, v_t, r_t>;

    std::vector<std::size_t> sizes_;
    std::size_t N_ = 0;
    std::vector<std::shared_ptr<detail::engine_box>> boxes_;

    std::pair<r_t, r_t> loss() {
        double s = 0, c = 0;
        for (auto &b : boxes_) {  // blocks; every engine has been launched by now
            double si = 0, ci = 0;
            CHECK_MGGCN(mggcn_engine_loss(b->e, &si, &ci));
            s += si;
            c += ci;
        }
        return {(r_t)(s / (double)N_), (r_t)(c / (double)N_)};
    }
    void inputs(const dist_row_dn_matrix<r_t> &H, const dist_row_dn_matrix<std::int32_t> &Y) {
        for (std::size_t i = 0; i < boxes_.size(); ++i) {
            auto &b = *boxes_[i];
            const bool nx = H[i].buffer() != b.x_seen, ny = Y[i].buffer() != b.y_seen;
            if (nx || ny)
                CHECK_MGGCN(mggcn_engine_set_inputs(b.e, nx ? H[i].buffer() : nullptr, ny ? Y[i].buffer() : nullptr, 0));
            b.x_seen = H[i].buffer();
            b.y_seen = Y[i].buffer();
        }
    }

   public:
    // Pass (A, A_T) exactly as src/main.cpp:151 does: forward multiplies by the blocks of A_T,
    // backward by the blocks of A (the reference swaps them for its layers, src/gcn.hpp:1023).
    // `sizes.back()` must already be padded to a multiple of P (src/main.cpp:135).
    dist_gcn(const dist_context &ctx, const csr_t &A, const csr_t &A_T, std::vector<std::size_t> sizes,
             bool residual_layer = false, gcn_options opt = {})
        : sizes_(std::move(sizes)), N_(A.n()) {
        if (residual_layer) throw std::invalid_argument("residual layers are not built (off in the reference CLI)");
        const int P = (int)ctx.size();
        std::vector<std::uint32_t> cuts(P + 1, 0);
        for (int i = 0; i < P; ++i) cuts[i + 1] = cuts[i] + (std::uint32_t)A[{(std::size_t)i, 0}].n();
        std::vector<std::int64_t> sz(sizes_.begin(), sizes_.end());
        for (int r = 0; r < P; ++r) {
            mggcn_engine_config cfg{};
            cfg.world = P;
            cfg.rank = r;
            cfg.device = (int)ctx[r].rank;
            cfg.num_layers = (int)sz.size() - 1;
            cfg.sizes = sz.data();
            cfg.n_global = (std::int64_t)N_;
            cfg.p = cuts.data();
            cfg.overlap = ctx.overlap ? 1 : 0;
            cfg.exact_grads = opt.exact_grads;
            cfg.use_graph = opt.use_graph && !opt.timers;
            cfg.timers = opt.timers;
            cfg.lrelu_slope = opt.lrelu_slope;
            cfg.comm_backend = MGGCN_COMM_P2P_GROUPED;
            auto box = std::make_shared<detail::engine_box>();
            CHECK_MGGCN(mggcn_engine_create(&cfg, &box->e));
            std::vector<block_t> fwd, bwd;
            for (int j = 0; j < P; ++j) {
                fwd.push_back(A_T[{(std::size_t)r, (std::size_t)j}]);
                bwd.push_back(A[{(std::size_t)r, (std::size_t)j}]);
            }
            detail::upload_block_row<block_t>(box->e, 0, fwd);
            detail::upload_block_row<block_t>(box->e, 1, bwd);
            boxes_.push_back(std::move(box));
        }
        if (P > 1) {
            std::size_t len = 0;
            CHECK_MGGCN(mggcn_engine_comm_handle(boxes_[0]->e, nullptr, 0, &len));
            std::vector<char> all(len * P);
            for (int r = 0; r < P; ++r) CHECK_MGGCN(mggcn_engine_comm_handle(boxes_[r]->e, all.data() + len * r, len, &len));
            for (int r = 0; r < P; ++r) CHECK_MGGCN(mggcn_engine_comm_connect(boxes_[r]->e, all.data(), len));
        }
    }

    // Every engine is launched before any of them is waited on: the stages of one GPU wait (on the
    // device) for blocks and barrier flags of the others.
    std::pair<r_t, r_t> train_forward(const dist_context &, const dist_row_dn_matrix<r_t> &H,
                                      const dist_row_dn_matrix<std::int32_t> &Y) {
        inputs(H, Y);
        for (auto &b : boxes_) CHECK_MGGCN(mggcn_engine_train_forward(b->e));
        return loss();
    }
    void backward(const dist_context &) {
        for (auto &b : boxes_) CHECK_MGGCN(mggcn_engine_backward(b->e));
    }
    void adam_update(const dist_context &, r_t lr, r_t beta1, r_t beta2, r_t wd, r_t eps) {
        for (auto &b : boxes_) CHECK_MGGCN(mggcn_engine_adam_update(b->e, lr, beta1, beta2, wd, eps));
    }
    std::pair<r_t, r_t> train_epoch(const dist_context &, const dist_row_dn_matrix<r_t> &H,
                                    const dist_row_dn_matrix<std::int32_t> &Y, r_t lr, r_t beta1, r_t beta2, r_t wd,
                                    r_t eps) {
        inputs(H, Y);
        // all graphs exist before the first one runs (see mggcn_engine_prepare in include/mggcn.h)
        for (auto &b : boxes_) CHECK_MGGCN(mggcn_engine_prepare(b->e, lr, beta1, beta2, wd, eps));
        for (auto &b : boxes_) CHECK_MGGCN(mggcn_engine_epoch(b->e, lr, beta1, beta2, wd, eps));
        return loss();
    }
    void update(const dist_context &, r_t, r_t) { throw std::logic_error("SGD update is not built; use adam_update"); }
    void sync() {
        for (auto &b : boxes_) CHECK_MGGCN(mggcn_engine_sync(b->e));
    }

    // replica 0's parameters and (reduced) gradients
    std::vector<gcn_layer_view> layers() const {
        std::vector<gcn_layer_view> out;
        for (std::size_t l = 0; l + 1 < sizes_.size(); ++l) out.emplace_back(boxes_[0]->e, (int)l, sizes_[l], sizes_[l + 1]);
        return out;
    }
    // "<prefix><gpu>_<name>:<ms>" like dist_context::dump_timers (src/dist_matrix.hpp:84-89)
    void dump_timers(std::ostream &out, const std::string &prefix) {
        for (std::size_t r = 0; r < boxes_.size(); ++r) {
            const std::string pre = prefix + std::to_string(r) + "_";
            std::size_t n = 0;
            CHECK_MGGCN(mggcn_engine_dump_timers(boxes_[r]->e, pre.c_str(), nullptr, 0, &n));
            std::string buf(n + 1, '\0');
            CHECK_MGGCN(mggcn_engine_dump_timers(boxes_[r]->e, pre.c_str(), buf.data(), buf.size(), &n));
            out.write(buf.data(), (std::streamsize)n);
        }
    }
    mggcn_engine_t *handle(std::size_t r) { return boxes_[r]->e; }
};

}  // namespace mg_gcn
