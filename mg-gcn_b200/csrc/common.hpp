// Shared helpers for the C-ABI implementation files (host + device).
#pragma once

#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>

#include "mggcn.h"

namespace mggcn {

// thread-local text behind mggcn_last_error()
char *last_error_buf();
int fail(int code, const char *fmt, ...) __attribute__((format(printf, 2, 3)));

extern std::atomic<uint64_t> g_launches;
inline void count_launch(uint64_t n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// splitmix64: counter-based, so generated data is independent of the thread count
static inline uint64_t mix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
static inline uint64_t hash2(uint64_t seed, uint64_t i) { return mix64(mix64(seed) ^ (i * 0xD6E8FEB86659FD93ull)); }
static inline double u01(uint64_t h) { return (double)(h >> 11) * (1.0 / 9007199254740992.0); }

// Description of the gathered multi-GPU operand of the SpMM (csrc/spmm.cu, used by csrc/engine.cu and
// mggcn_spmm_gathered_f32_i32): see spmm_args::g.
struct spmm_gather {
    const uint32_t *flags = nullptr;  // P words; flags[s] != 0: the block of stage s is in place (flags[0] unused)
    const float *B_own = nullptr;     // this rank's own rows (stage 0), leading dimension = ldb; copied into the
                                      // gather buffer ahead of the launch (null: they are there already)
    uint32_t own_lo = 0, own_n = 0;   // global column range of the own rows
    uint32_t ncols = 0;               // columns of the gathered operand (set by the dispatcher)
    int P = 1, rank = 0;
    int *err = nullptr;               // raised to 2 when a stage flag never arrives
    uint32_t cuts[16] = {0};          // cuts[i] = first global column of block i (i < P), 0xFFFFFFFF beyond
};

}  // namespace mggcn

#ifdef __CUDACC__
#include <cuda_runtime.h>
#define MGGCN_CUDA_TRY(expr)                                                             \
    do {                                                                                 \
        cudaError_t e__ = (expr);                                                        \
        if (e__ != cudaSuccess)                                                          \
            return ::mggcn::fail((int)e__, "%s: %s", #expr, cudaGetErrorString(e__));   \
    } while (0)
#define MGGCN_LAUNCH_CHECK()                                                             \
    do {                                                                                 \
        cudaError_t e__ = cudaGetLastError();                                            \
        if (e__ != cudaSuccess)                                                          \
            return ::mggcn::fail((int)e__, "kernel launch (%s:%d): %s", __FILE__, __LINE__, \
                                 cudaGetErrorString(e__));                               \
        ::mggcn::count_launch();                                                         \
    } while (0)
#endif
