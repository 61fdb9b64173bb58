// matrix.hpp -- per-GPU runtime (`context`) and the two containers of the hot path
// (`csr_matrix`, `dn_matrix`).  Same public surface as the reference's src/matrix.hpp:39-639
// (SURVEY.md Appendix E); the bodies sit on the C-ABI of include/mggcn.h.
//
// Differences that a caller can observe:
//   * no cuSPARSE / cuBLAS handles and no cuSPARSE descriptors: `get_mat()` is gone, the SpMM takes
//     raw pointers (a plan -- the degree-aware work list of csrc/spmm.cu -- is cached per matrix);
//   * events are interned: `record` / `wait` look a name up once and reuse the handle;
//   * `normalize` and `transpose` are the reference's serial forms (deterministic, sorted rows),
//     run multi-threaded through the ABI (src/matrix.hpp:377-388, :426-453).
#pragma once

#include <algorithm>
#include <chrono>
#include <filesystem>
#include <iostream>
#include <map>
#include <mutex>
#include <ostream>
#include <tuple>
#include <unordered_map>
#include <utility>
#include <vector>

#include "mg_gcn.hpp"

namespace mg_gcn {

// -------------------------------------------------------------------------------------------------
// context: one GPU.  Stream 0 = compute (low priority), stream 1 = exchange (high priority), named
// events and timers with the reference's CSV format "<prefix><name>:<ms>\n" (src/matrix.hpp:69-158).
// A cheap value type: copies share streams, events and timers.
// -------------------------------------------------------------------------------------------------
class context {
   public:
    using event_t = std::shared_ptr<CUevent_st>;
    using stream_t = std::shared_ptr<CUstream_st>;
    using events_t = std::shared_ptr<std::unordered_map<std::string, event_t>>;
    using timers_t = std::shared_ptr<std::map<std::string, std::pair<event_t, event_t>>>;
    using walltimes_t = std::shared_ptr<std::map<std::string, std::chrono::time_point<std::chrono::steady_clock>>>;

    std::size_t rank = 0;
    std::vector<stream_t> cuda_streams;
    events_t events;
    timers_t timers;
    walltimes_t walltimes;

    context() = default;
    explicit context(std::size_t device)
        : rank(device),
          events(std::make_shared<events_t::element_type>()),
          timers(std::make_shared<timers_t::element_type>()),
          walltimes(std::make_shared<walltimes_t::element_type>()) {
        set();
        int lo = 0, hi = 0;
        CHECK_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        for (int prio : {lo, hi}) {
            cudaStream_t s = nullptr;
            CHECK_CUDA(cudaStreamCreateWithPriority(&s, cudaStreamNonBlocking, prio));
            const int dev = (int)device;
            cuda_streams.emplace_back(s, [dev](cudaStream_t q) {
                cudaSetDevice(dev);
                cudaStreamDestroy(q);
            });
        }
    }

    void set() const { CHECK_CUDA(cudaSetDevice((int)rank)); }
    void sync() const {
        set();
        for (auto &s : cuda_streams) CHECK_CUDA(cudaStreamSynchronize(s.get()));
    }
    cudaStream_t stream(std::size_t id = 0) const { return cuda_streams[id].get(); }

    event_t create_event() const {
        set();
        cudaEvent_t e = nullptr;
        CHECK_CUDA(cudaEventCreate(&e));
        const int dev = (int)rank;
        return event_t(e, [dev](cudaEvent_t q) {
            cudaSetDevice(dev);
            cudaEventDestroy(q);
        });
    }
    stream_t create_stream() const {
        set();
        cudaStream_t s = nullptr;
        CHECK_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
        return stream_t(s, [](cudaStream_t q) { cudaStreamDestroy(q); });
    }

    void record(const std::string &name, std::size_t stream_id) {
        event_t &e = (*events)[name];
        if (!e) e = create_event();
        set();
        CHECK_CUDA(cudaEventRecord(e.get(), stream(stream_id)));
    }
    void wait(const std::string &name, std::size_t stream_id) {
        auto it = events->find(name);
        if (it == events->end()) throw std::logic_error("context::wait on an event never recorded: " + name);
        set();
        CHECK_CUDA(cudaStreamWaitEvent(stream(stream_id), it->second.get(), 0));
    }
    void register_timer(const std::string &name, const std::string &beg, const std::string &end) {
        (*timers)[name] = {(*events)[beg], (*events)[end]};
    }
    void register_walltime(const std::string &name, std::size_t stream_id) {
        auto &slot = (*walltimes)[name];
        set();
        CHECK_CUDA(cudaLaunchHostFunc(stream(stream_id), [](void *p) {
            *static_cast<std::chrono::time_point<std::chrono::steady_clock> *>(p) = std::chrono::steady_clock::now();
        }, &slot));
    }
    long long measure_walltime(const std::string &name) {
        return std::chrono::duration_cast<std::chrono::milliseconds>((*walltimes)[name].time_since_epoch()).count();
    }
    static float measure(const timers_t &timers, const std::string &name) {
        float ms = 0.f;
        auto it = timers->find(name);
        if (it != timers->end() && it->second.first && it->second.second)
            CHECK_CUDA(cudaEventElapsedTime(&ms, it->second.first.get(), it->second.second.get()));
        return ms;
    }
    float measure(const std::string &name) const { return measure(timers, name); }
    static void dump_timers(const timers_t &timers, std::ostream &out, const std::string &prefix) {
        for (auto &t : *timers) out << prefix << t.first << ':' << measure(timers, t.first) << '\n';
    }
    void dump_timers(std::ostream &out, const std::string &prefix) { dump_timers(timers, out, prefix); }
};

template <typename r_t>
class dn_matrix;

}  // namespace mg_gcn

#include "matrix_dn.hpp"
#include "matrix_csr.hpp"
