// mg_gcn.hpp -- conventions shared by the C++ surface: error policy, buffer ownership, types.
//
// Mirrors the role of the reference's src/mg_gcn.hpp:1-132 (CHECK_* macros that print and exit,
// cuda_ptr<T> = shared_ptr<T[]> with a CUDA deleter, cuda_malloc / cuda_malloc_managed).  The
// surface is header-only C++17 over the C-ABI of include/mggcn.h; it needs the CUDA runtime
// headers only for allocation, copies, streams and events -- every kernel lives behind the ABI.
#pragma once

#include <cuda_runtime_api.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>

#include "mggcn.h"

namespace mg_gcn {

// Fail-fast policy of the reference's C++ surface (src/mg_gcn.hpp:31-68): print and exit.
// The C-ABI below never exits; it hands back a code and a message.
#define CHECK_CUDA(call)                                                                           \
    do {                                                                                           \
        cudaError_t status__ = (call);                                                             \
        if (status__ != cudaSuccess) {                                                             \
            std::printf("CUDA failure at %s:%d: %s (%d)\n", __FILE__, __LINE__,                    \
                        cudaGetErrorString(status__), (int)status__);                              \
            std::exit(EXIT_FAILURE);                                                               \
        }                                                                                          \
    } while (0)

#define CHECK_MGGCN(call)                                                                          \
    do {                                                                                           \
        int status__ = (call);                                                                     \
        if (status__ != MGGCN_OK) {                                                                \
            std::printf("mggcn failure at %s:%d: %s (%d)\n", __FILE__, __LINE__,                   \
                        mggcn_last_error(), status__);                                             \
            std::exit(EXIT_FAILURE);                                                               \
        }                                                                                          \
    } while (0)

// bad files / bad arguments are exceptions, caught in main (src/matrix.hpp:33-37, src/main.cpp:198-209)
class matrix_error : public std::runtime_error {
   public:
    template <typename T>
    explicit matrix_error(T t) : std::runtime_error(t) {}
};

template <typename T>
using cuda_ptr = std::shared_ptr<T[]>;

// True when a CUDA device can be used by this process.  Host-only tools (file conversion,
// pre-processing, the generator) run without one; every compute entry point needs one.
inline bool have_device() {
    static const bool ok = [] {
        int n = 0;
        const cudaError_t e = cudaGetDeviceCount(&n);
        if (e != cudaSuccess) cudaGetLastError();
        return e == cudaSuccess && n > 0;
    }();
    return ok;
}

// Host-indexable buffers, like the reference's cudaMallocManaged containers (src/mg_gcn.hpp:74-82):
// callers index matrices on the host (operator[]) and hand the same pointer to kernels.  Without a
// device the allocation is plain host memory so that the host half of the surface still works.
// MGGCN_POISON=1 fills every new buffer with 0xFF bytes (NaN as float): a read of memory nobody wrote shows up as
// NaN instead of depending on what the previous process left in device memory (debugging aid of the surface tests)
inline bool poison_allocations() {
    static const bool on = std::getenv("MGGCN_POISON") != nullptr;
    return on;
}

template <typename T>
cuda_ptr<T> cuda_malloc_managed(std::size_t count) {
    if (count == 0) count = 1;
    T *p = nullptr;
    if (have_device()) {
        CHECK_CUDA(cudaMallocManaged(reinterpret_cast<void **>(&p), count * sizeof(T)));
        if (poison_allocations()) CHECK_CUDA(cudaMemset(p, 0xFF, count * sizeof(T)));
        return cuda_ptr<T>(p, [](T *q) { cudaFree(q); });
    }
    p = static_cast<T *>(std::malloc(count * sizeof(T)));
    if (!p) throw std::bad_alloc();
    return cuda_ptr<T>(p, [](T *q) { std::free(q); });
}

// Device-only buffers (src/mg_gcn.hpp:84-90).
template <typename T>
cuda_ptr<T> cuda_malloc(std::size_t count) {
    if (count == 0) count = 1;
    T *p = nullptr;
    CHECK_CUDA(cudaMalloc(reinterpret_cast<void **>(&p), count * sizeof(T)));
    if (poison_allocations()) CHECK_CUDA(cudaMemset(p, 0xFF, count * sizeof(T)));
    return cuda_ptr<T>(p, [](T *q) { cudaFree(q); });
}

// Moves a managed range next to the GPU that is about to read it, so the first epoch does not pay
// page faults (the reference never prefetches or advises).
inline void prefetch_to_device(const void *p, std::size_t bytes, int device, cudaStream_t st = nullptr) {
    if (!have_device() || !p || !bytes) return;
    cudaPointerAttributes a{};
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess || a.type != cudaMemoryTypeManaged) {
        cudaGetLastError();
        return;
    }
#if CUDART_VERSION >= 13000
    cudaMemLocation loc{cudaMemLocationTypeDevice, device};
    cudaMemPrefetchAsync(p, bytes, loc, 0, st);
#else
    cudaMemPrefetchAsync(p, bytes, device, st);
#endif
    // The migration is finished before anything else is queued: the callers are constructors (scatter of a dataset,
    // block split), and the first consumers of these ranges are copy-engine pulls from ANOTHER device and TMA loads,
    // which should not have to race a migration still in flight on the default stream.
    cudaStreamSynchronize(st);
    cudaGetLastError();
}

// Keeps a managed range where it is: preferred location = `device`, and a direct mapping for every other GPU and the
// host, so that a peer copy, a kernel on another GPU or a host read reaches the pages over NVLink / PCIe instead of
// migrating them.  The multi-GPU containers (one block or one replica per GPU, read by peers in every exchange stage
// and every all-reduce) call it for each block; the reference leaves placement to the driver (src/mg_gcn.hpp:74-82).
// MGGCN_PIN_MANAGED=0 switches it off, =dist / =repl keeps it for one container family only (A/B runs of the
// layer-stack stress).  `family`: 'd' = dist_row_dn_matrix block, 'r' = repl_dn_matrix replica.
inline void pin_managed(const void *p, std::size_t bytes, int device, char family = 'd') {
    static const std::string mode = std::getenv("MGGCN_PIN_MANAGED") ? std::getenv("MGGCN_PIN_MANAGED") : "all";
    const bool on = mode == "all" || (mode == "dist" && family == 'd') || (mode == "repl" && family == 'r');
    if (!on || !have_device() || !p || !bytes) return;
    cudaPointerAttributes a{};
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess || a.type != cudaMemoryTypeManaged) {
        cudaGetLastError();
        return;
    }
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) n = 0;
#if CUDART_VERSION >= 13000
    auto advise = [&](cudaMemoryAdvise what, int dev) {
        cudaMemLocation loc{dev == cudaCpuDeviceId ? cudaMemLocationTypeHost : cudaMemLocationTypeDevice, dev == cudaCpuDeviceId ? 0 : dev};
        cudaMemAdvise(p, bytes, what, loc);
    };
#else
    auto advise = [&](cudaMemoryAdvise what, int dev) { cudaMemAdvise(p, bytes, what, dev); };
#endif
    advise(cudaMemAdviseSetPreferredLocation, device);
    for (int d = 0; d < n; ++d) {
        int can = d == device;
        if (!can) cudaDeviceCanAccessPeer(&can, d, device);
        if (can) advise(cudaMemAdviseSetAccessedBy, d);
    }
    advise(cudaMemAdviseSetAccessedBy, cudaCpuDeviceId);
    cudaGetLastError();
}

}  // namespace mg_gcn
