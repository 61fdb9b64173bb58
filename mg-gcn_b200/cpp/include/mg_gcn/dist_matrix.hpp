// dist_matrix.hpp -- multi-GPU runtime and the row-partitioned containers of the hot path:
// dist_context, dist_row_csr_matrix (P x P blocks), dist_row_dn_matrix (row blocks + staged
// broadcast), repl_dn_matrix (replicated weights + all-reduce).  Surface of the reference's
// src/dist_matrix.hpp:12-90,170-260,394-639.  The dead column-partition classes are not provided
// (DESIGN.md section 0).
//
// One process drives all P GPUs, like the reference.  The exchange is not NCCL here: every GPU
// pair is peer-mapped over NVLink and blocks move with copy-engine peer copies ordered by
// cross-device events, so a broadcast is P concurrent pulls from the root.
#pragma once

#include "matrix.hpp"

namespace mg_gcn {

class dist_context {
    std::vector<context> ctxs_;

   public:
    const bool overlap;

    dist_context(std::size_t P, bool overlap = true) : overlap(overlap) {
        for (std::size_t i = 0; i < P; ++i) ctxs_.emplace_back(i);
        for (std::size_t i = 0; i < P; ++i) {
            ctxs_[i].set();
            for (std::size_t j = 0; j < P; ++j) {
                if (i == j) continue;
                int can = 0;
                CHECK_CUDA(cudaDeviceCanAccessPeer(&can, (int)i, (int)j));
                if (!can) continue;  // copies then stage through the host; still correct
                const cudaError_t e = cudaDeviceEnablePeerAccess((int)j, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CHECK_CUDA(e);
                cudaGetLastError();
            }
        }
    }

    // the stream the exchange runs on: 1 = dedicated high-priority stream, 0 = the compute stream (-S)
    int bcast_stream_id() const { return overlap ? 1 : 0; }
    std::size_t size() const { return ctxs_.size(); }
    void sync() const {
        for (auto &c : ctxs_) c.sync();
    }
    context operator[](std::size_t i) const { return ctxs_[i]; }
    // The reference hands out ncclComm_t here (src/dist_matrix.hpp:45-47); this build has no
    // communicator object -- peers are addressed by device ordinal.
    void *operator()(std::size_t) const { return nullptr; }

    void record(const std::string &name, std::size_t stream_id) {
        for (auto &c : ctxs_) c.record(name, stream_id);
    }
    void wait(const std::string &name, std::size_t stream_id) {
        for (auto &c : ctxs_) c.wait(name, stream_id);
    }
    void register_timer(const std::string &name, const std::string &beg, const std::string &end) {
        for (auto &c : ctxs_) c.register_timer(name, beg, end);
    }
    void register_walltime(const std::string &name, std::size_t stream_id) {
        for (auto &c : ctxs_) c.register_walltime(name, stream_id);
    }
    long long measure_walltime(const std::string &name) {
        long long v = 0;
        for (auto &c : ctxs_) v = std::max(v, c.measure_walltime(name));
        return v;
    }
    float measure(const std::string &name) const {
        float v = 0.f;
        for (auto &c : ctxs_) v = std::max(v, c.measure(name));
        return v;
    }
    // "<prefix><gpu>_<name>:<ms>" per GPU (src/dist_matrix.hpp:84-89)
    void dump_timers(std::ostream &out, const std::string &prefix = "") {
        for (std::size_t i = 0; i < ctxs_.size(); ++i) ctxs_[i].dump_timers(out, prefix + std::to_string(i) + "_");
    }
};

// -------------------------------------------------------------------------------------------------
// dist_row_csr_matrix: As[i][j] = rows p[i]..p[i+1], columns q[j]..q[j+1] of A with block-local
// column ids, within-row order preserved (src/dist_matrix.hpp:215-259).  Bit-exact with the
// reference's extraction (tests/test_host.py).  Block row i lives next to GPU i.
// -------------------------------------------------------------------------------------------------
template <typename x_t, typename v_t, typename r_t>
class dist_row_csr_matrix {
    using matrix_t = csr_matrix<x_t, v_t, r_t>;
    std::vector<std::vector<matrix_t>> As;
    v_t N_ = 0, M_ = 0;

   public:
    dist_row_csr_matrix() = default;
    dist_row_csr_matrix(const dist_context &ctx, const matrix_t &A, const std::vector<v_t> &p, const std::vector<v_t> &q)
        : N_(A.n()), M_(A.m()) {
        const int P = (int)p.size() - 1, Q = (int)q.size() - 1;
        std::vector<std::uint64_t> w((std::size_t)A.n() + 1);
        for (std::size_t i = 0; i <= A.n(); ++i) w[i] = (std::uint64_t)A.indptr()[i];
        std::vector<std::vector<std::uint64_t>> rp((std::size_t)P * Q);
        std::vector<std::uint64_t *> rpp(rp.size());
        for (int i = 0; i < P; ++i)
            for (int j = 0; j < Q; ++j) {
                rp[(std::size_t)i * Q + j].assign((std::size_t)(p[i + 1] - p[i]) + 1, 0);
                rpp[(std::size_t)i * Q + j] = rp[(std::size_t)i * Q + j].data();
            }
        const auto *pc = reinterpret_cast<const std::uint32_t *>(p.data());
        const auto *qc = reinterpret_cast<const std::uint32_t *>(q.data());
        const auto *ci = reinterpret_cast<const std::uint32_t *>(A.indices());
        CHECK_MGGCN(mggcn_csr_block_counts(A.n(), w.data(), ci, P, pc, Q, qc, rpp.data()));
        std::vector<cuda_ptr<v_t>> bi(rp.size());
        std::vector<cuda_ptr<r_t>> bd(rp.size());
        std::vector<std::uint32_t *> bip(rp.size());
        std::vector<float *> bdp(rp.size());
        for (std::size_t k = 0; k < rp.size(); ++k) {
            bi[k] = cuda_malloc_managed<v_t>((std::size_t)rp[k].back());
            bd[k] = cuda_malloc_managed<r_t>((std::size_t)rp[k].back());
            bip[k] = reinterpret_cast<std::uint32_t *>(bi[k].get());
            bdp[k] = bd[k].get();
        }
        CHECK_MGGCN(mggcn_csr_block_fill(A.n(), w.data(), ci, A.data(), P, pc, Q, qc, rpp.data(), bip.data(), bdp.data()));
        As.resize(P);
        for (int i = 0; i < P; ++i)
            for (int j = 0; j < Q; ++j) {
                const std::size_t k = (std::size_t)i * Q + j;
                auto bp = cuda_malloc_managed<x_t>(rp[k].size());
                for (std::size_t r = 0; r < rp[k].size(); ++r) bp[r] = (x_t)rp[k][r];
                const int dev = i < (int)ctx.size() ? i : 0;
                prefetch_to_device(bp.get(), rp[k].size() * sizeof(x_t), dev);
                prefetch_to_device(bi[k].get(), (std::size_t)rp[k].back() * sizeof(v_t), dev);
                prefetch_to_device(bd[k].get(), (std::size_t)rp[k].back() * sizeof(r_t), dev);
                As[i].emplace_back(bp, bi[k], bd[k], (v_t)(p[i + 1] - p[i]), (v_t)(q[j + 1] - q[j]));
            }
    }

    v_t n() const { return N_; }
    v_t m() const { return M_; }
    std::size_t size() const { return As.size(); }
    std::size_t size(std::size_t i) const { return As[i].size(); }
    matrix_t operator[](std::pair<std::size_t, std::size_t> ij) const { return As[ij.first][ij.second]; }
};

}  // namespace mg_gcn

#include "dist_matrix_dn.hpp"
