// Epoch engine: one handle per rank (= per GPU).  Mirrors gcn / dist_gcn<true,...>
// (src/gcn.hpp:937-1056), gcn_layer / dist_gcn_layer (:411-637), sparse_linear / dist_sparse_linear
// (:13-86), dist_row_linear (:191-296), the loss (:872-935) and the staged SpMM with its double
// buffered exchange (src/cuda_utils.hpp:57-92, src/dist_matrix.hpp:458-467).
//
// What is different from the reference, by design:
//   * device-resident buffers (one arena per rank) instead of cudaMallocManaged;
//   * activation / activation-gradient / bias are epilogues of the kernel that produces the tensor
//     (SpMM or GEMM), so every activation matrix makes one HBM round trip;
//   * the exchange of the feature blocks has two back-ends: "nccl" = the reference's root-rotating
//     ncclBroadcast schedule, and "p2p" = every rank pulls block (rank+s)%P at stage s straight out
//     of the owner's arena over NVLink (copy engine, peer-mapped memory), so each stage is a
//     permutation and every link is busy in every stage; stage 0 is the local block and needs no
//     transfer at all.  Cross-rank ordering uses a flag barrier in peer-mapped memory;
//   * weight gradients of all layers are reduced in one step (one ncclAllReduce, or in p2p mode a
//     kernel that sums the peers' packed gradients in rank order) and Adam is one launch;
//   * the whole epoch is captured into one CUDA graph (use_graph), removing ~100 launches and all
//     the event bookkeeping of src/cuda_utils.hpp:61-91 from the host path.
#include <cuda.h>  // CUstream / CUdeviceptr for cuStreamWriteValue32 (resolved at run time, no -lcuda)
#include <dlfcn.h>
#include <nccl.h>
#include <unistd.h>

#include <algorithm>
#include <cmath>
#include <string>
#include <vector>

#include "common.hpp"

using namespace mggcn;

namespace mggcn {
// mggcn_tune("skip_exchange", 1): the peer copies of the p2p back-ends are not enqueued (barriers, events and
// launches stay).  The result is wrong by construction -- it exists to MEASURE how much of an epoch the exchange
// leaves exposed: exposed = epoch(with copies) - epoch(without).  Never set outside a measurement.
int g_skip_exchange = 0;
// exchange tuning: a block's pull is cut into g_xstreams row ranges issued on as many streams (one copy engine each);
// g_xpull = 1 moves the bytes with pull_copy_kernel (SM loads over NVLink) instead of the copy engines
int g_xstreams = 0;  // 0 = by the block size (see peer_pull)
int g_xpull = 0;
// > 0: set_inputs reads pinned host rows with that many CTAs of pull_copy_kernel instead of a copy-engine transfer
int g_h2d_sm = 0;
// > 0: a contiguous host -> device input transfer is issued in pieces of this many KB (see copy_rows_in)
int g_input_flip = 1;  // single-GPU engines swap pointers instead of copying prefetched inputs (swap_in_inputs)
int g_h2d_chunk_kb = 0;  // 0 = by size on multi-GPU engines (h2d_piece_bytes); a value past the transfer size = one transfer
// grouping of the stages of the grouped exchange: 0 = by the shape (see choose_groups), 1 = one launch per stage,
// 2 = own | 1 | 2-3 | 4-7 | ...
int g_xgroups = 0;
void spmm_touched_cols_hint(int64_t n);
int spmm_gathered(cudaStream_t st, int64_t rows, int64_t cols, const uint32_t *rowptr, const uint32_t *colidx,
                  const float *vals, float *gathered, int64_t ld, float *C, int64_t ldc, int F, float alpha,
                  float beta, const mggcn_epilogue *epi, const mggcn_spmm_plan_t *plan, void *ws, const spmm_gather &g);
}

namespace {

// ---------------------------------------------------------------------------------------------
// NCCL is bound at run time (dlopen), so the library loads without it for P = 1 and shares the
// copy torch already mapped when the host program is Python.
// ---------------------------------------------------------------------------------------------
struct nccl_api {
    void *h = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
nccl_api g_nccl;

int nccl_load() {
    if (g_nccl.h) return MGGCN_OK;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    void *h = nullptr;
    for (const char *n : names) {
        h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) return fail(MGGCN_E_NCCL, "cannot load libnccl: %s", dlerror());
#define BIND(field, sym)                                                        \
    g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(h, sym));     \
    if (!g_nccl.field) return fail(MGGCN_E_NCCL, "libnccl lacks %s", sym)
    BIND(GetUniqueId, "ncclGetUniqueId");
    BIND(CommInitRank, "ncclCommInitRank");
    BIND(CommDestroy, "ncclCommDestroy");
    BIND(Broadcast, "ncclBroadcast");
    BIND(AllReduce, "ncclAllReduce");
    BIND(GroupStart, "ncclGroupStart");
    BIND(GroupEnd, "ncclGroupEnd");
    BIND(GetErrorString, "ncclGetErrorString");
#undef BIND
    g_nccl.h = h;
    return MGGCN_OK;
}

#define NCCL_TRY(expr)                                                                          \
    do {                                                                                        \
        ncclResult_t r__ = (expr);                                                              \
        if (r__ != ncclSuccess)                                                                 \
            return fail(MGGCN_E_NCCL, "%s: %s", #expr, g_nccl.GetErrorString(r__));             \
    } while (0)
#define TRY(expr)                  \
    do {                           \
        int rc__ = (expr);         \
        if (rc__) return rc__;     \
    } while (0)

inline int64_t up4(int64_t x) { return (x + 3) / 4 * 4; }

// MGGCN_TRACE=1: host-side time stamps of the enqueue path (finds calls that block the launching thread)
inline bool trace_on() {
    static const bool on = getenv("MGGCN_TRACE") != nullptr;
    return on;
}
inline double trace_ms() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}
#define TRACE(rank, label)                                                               \
    do {                                                                                 \
        if (trace_on()) fprintf(stderr, "[trace r%d] %.3f %s\n", (rank), trace_ms(), (label)); \
    } while (0)
inline size_t up256(size_t x) { return (x + 255) / 256 * 256; }

// ---------------------------------------------------------------------------------------------
// device helpers of the p2p back-end
// ---------------------------------------------------------------------------------------------
constexpr int kMaxRanks = 16;
constexpr int kBarrierSlots = 128;
struct peer_table {
    uint32_t *flags[kMaxRanks];  // peer r's flag page (kBarrierSlots x kMaxRanks uint32)
};

// Flag barrier over peer-mapped memory.  Slot k has its own monotonically increasing ticket, kept
// on the device so that a captured graph can be replayed.  Every rank runs the same sequence of
// barriers, so the tickets agree.  A timeout (about 10 s of SM clock) raises *err instead of hanging.
__global__ void xbarrier_kernel(peer_table pt, uint32_t *tickets, int slot, int rank, int P, int *err,
                                long long timeout_cycles) {
    __shared__ uint32_t ticket;
    if (threadIdx.x == 0) ticket = ++tickets[slot];
    __syncthreads();
    const int t = threadIdx.x;
    if (t < P) {
        __threadfence_system();
        volatile uint32_t *dst = pt.flags[t] + slot * kMaxRanks + rank;
        *dst = ticket;
        __threadfence_system();
        volatile uint32_t *src = pt.flags[rank] + slot * kMaxRanks + t;
        const long long t0 = clock64();
        while ((int32_t)(*src - ticket) < 0) {
            if (clock64() - t0 > timeout_cycles) {
                if (atomicCAS(err, 0, 1) == 0) {  // first failure wins: which flag was missing
                    err[1] = slot * kMaxRanks + t;
                    err[2] = (int)*src;
                    err[3] = (int)ticket;
                }
                break;
            }
            __nanosleep(200);
        }
        __threadfence_system();
    }
}

// SM-driven pull of a peer's feature block over NVLink (128-bit loads from peer-mapped memory).
// Used instead of a copy-engine transfer when the peer GPU is driven by THIS process and the epoch
// is enqueued eagerly: there a cross-device cudaMemcpyAsync waiting behind this GPU's stage barrier
// holds back the peer's own launches (measured on B200: the peer's barrier kernel did not start
// until the waiting one timed out), which a kernel that touches only this GPU's queues cannot do.
__global__ void __launch_bounds__(512) pull_copy_kernel(const float4 *__restrict__ src, float4 *__restrict__ dst, size_t n4) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n4; i += 4 * stride) {
        const float4 a = src[i], b = src[i + stride], c = src[i + 2 * stride], d = src[i + 3 * stride];
        dst[i] = a; dst[i + stride] = b; dst[i + 2 * stride] = c; dst[i + 3 * stride] = d;
    }
    for (; i < n4; i += stride) dst[i] = src[i];
}

// Publishes a stage flag behind the transfers queued on `st`.  A stream memory operation is executed by the
// GPU's front end: unlike a small memset (which may be a kernel) it needs no SM slot, so it cannot be starved by
// a grid of CTAs that wait for this very flag.
typedef CUresult (*stream_write32_fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
stream_write32_fn stream_write32() {
    static stream_write32_fn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (getenv("MGGCN_FLAG_MEMSET") || cudaGetDriverEntryPoint("cuStreamWriteValue32", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            p = nullptr;
        }
        return reinterpret_cast<stream_write32_fn>(p);
    }();
    return fn;
}
bool g_memop_in_capture = true;  // cleared when a capture that used stream memory operations fails
int publish_flag(cudaStream_t st, uint32_t *flag, bool capturing) {
    stream_write32_fn w = capturing && !g_memop_in_capture ? nullptr : stream_write32();
    if (w) {
        const CUresult r = w(reinterpret_cast<CUstream>(st), reinterpret_cast<CUdeviceptr>(flag), 1u, 0u);
        if (r == CUDA_SUCCESS) return MGGCN_OK;
        if (r != CUDA_ERROR_NOT_SUPPORTED && r != CUDA_ERROR_STREAM_CAPTURE_UNSUPPORTED)
            return fail((int)r, "cuStreamWriteValue32 failed (%d)", (int)r);
    }
    MGGCN_CUDA_TRY(cudaMemsetAsync(flag, 0x01, sizeof(uint32_t), st));
    return MGGCN_OK;
}

__global__ void pad_rows_kernel(const float *src, int64_t rows, int64_t cols, float *dst, int64_t ld) {
    const int64_t total = rows * ld;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / ld, c = i - r * ld;
        dst[i] = c < cols ? src[r * cols + c] : 0.f;
    }
}

struct grad_peers {
    const float *g[kMaxRanks];
};
// sum of the peers' packed gradients in rank order (identical on every rank)
__global__ void grad_reduce_kernel(grad_peers gp, int P, size_t n, float *out) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int r = 0; r < P; ++r) s += gp.g[r][i];
        out[i] = s;
    }
}

// step counter and bias corrections live on the device so the Adam launch can sit in a graph
// bc1 = (float)(1 - pow(beta1, t)), src/gcn.hpp:256-257
__global__ void adam_prep_kernel(float *state, float beta1, float beta2) {
    const float t = state[0] + 1.f;
    state[0] = t;
    state[1] = (float)(1.0 - pow((double)beta1, (double)t));
    state[2] = (float)(1.0 - pow((double)beta2, (double)t));
}
__global__ void adam_dev_kernel(size_t n, float *p, float *g, float *m, float *v, const float *state, float lr,
                                float beta1, float beta2, float wd, float eps, size_t n_decay, const int *err) {
    // a peer that never arrived (barrier / stage-flag timeout) leaves partial sums behind: keep the parameters
    // as they are, the host sees the error at its next loss() / sync()
    if (err && err[0] != 0) return;
    const float bc1 = state[1], bc2 = state[2];
    const float step_size = lr / bc1;
    const float a1 = 1.f - beta1, a2 = 1.f - beta2;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        float gi = g[i];
        if (i < n_decay) {
            gi = fmaf(wd, p[i], gi);
            g[i] = gi;
        }
        const float mi = a1 * gi + beta1 * m[i];
        const float vi = a2 * gi * gi + beta2 * v[i];
        m[i] = mi;
        v[i] = vi;
        p[i] -= __fdividef(step_size * mi, sqrtf(__fdividef(vi, bc2)) + eps);
    }
}

struct adj_block {
    uint32_t *rowptr = nullptr, *colidx = nullptr;
    float *vals = nullptr;
    int64_t rows = 0, cols = 0, nnz = 0;
    mggcn_spmm_plan_t *plan = nullptr;
};

struct layer_t {
    int64_t in = 0, out = 0;
    bool activation = false, gemm_first = false;
    float *W = nullptr, *b = nullptr, *GW = nullptr, *Gb = nullptr;  // views into the packed vectors
    // the tensor-core GEMM reads W through TMA (16-byte aligned base and row pitch): when out % 4 != 0 (41
    // classes) the forward pass refreshes a copy with rows padded to a multiple of 4 floats and the GEMMs of the
    // epoch read that one; otherwise Wt == W
    float *Wt = nullptr, *Wpad = nullptr;
    int64_t ldw = 0;
    float *Z = nullptr;   // layer output, n x up4(out)
    float *AH = nullptr;  // exact_grads && !gemm_first: saved A*H, n x up4(in)
};

struct named_timer {
    std::string name;
    cudaEvent_t beg, end;
    std::string cfg;  // SpMM regions: the kernel instance and schedule that ran inside (mggcn_last_spmm_config)
};

}  // namespace

struct mggcn_engine {
    int P = 1, rank = 0, device = 0, L = 0;
    bool overlap = true, exact = false, use_graph = false, p2p = false, timers_on = false;
    float slope = 0.01f;
    int64_t N = 0, n = 0;  // global / local rows
    std::vector<uint32_t> cuts;
    std::vector<int64_t> sizes;
    std::vector<layer_t> layers;
    cudaStream_t sc = nullptr, sm = nullptr, scopy = nullptr;
    static constexpr int kXStreams = 4;
    cudaStream_t sx[kXStreams] = {nullptr};      // extra exchange streams: sx[0] is sm itself
    cudaEvent_t ev_xfork = nullptr, ev_xjoin[kXStreams] = {nullptr};
    ncclComm_t comm = nullptr;

    adj_block adj[2][kMaxRanks];
    bool adj_set[2] = {false, false};

    // arena
    char *arena = nullptr;
    size_t arena_bytes = 0;
    uint32_t *flags = nullptr;      // kBarrierSlots x kMaxRanks (first page of the arena)
    float *X = nullptr;
    int32_t *Y = nullptr;
    int64_t ldX = 0;
    float *HW = nullptr, *G[2] = {nullptr, nullptr}, *stage[2] = {nullptr, nullptr};
    // gathered form (p2p back-end): one launch per aggregation over the whole block row; the copy engines
    // fill `gather` (all ranks' rows, global order) while the kernel already works on the stages that landed
    bool fused = false;
    // grouped form: same all-gather buffer, but plain launches over GROUPS of stages -- own block | stage 1 |
    // stages 2-3 | 4-7 | 8-15 -- each ordered behind the transfer of its last block with a stream event
    bool grouped = false;
    std::vector<std::pair<int, int>> groups;  // [first stage, end stage) per launch
    bool groups_chosen = false;               // fixed by the first adjacency operand (choose_groups)
    float *gather = nullptr;
    uint32_t *stage_flags = nullptr;
    cudaEvent_t ev_bar = nullptr, ev_copied = nullptr;
    int64_t maxw = 0, gw = 0, max_block_rows = 0;
    float *params = nullptr, *grads = nullptr, *gsum = nullptr, *adam_m = nullptr, *adam_v = nullptr;
    size_t n_params = 0, n_weights = 0;
    float *adam_state = nullptr;  // t, bc1, bc2
    double *loss_out = nullptr;   // sum |log p_y|, #correct
    void *ws_gemm = nullptr, *ws_colsum = nullptr, *ws_loss = nullptr, *ws_spmm = nullptr;
    size_t ws_gemm_bytes = 0, ws_colsum_bytes = 0, ws_loss_bytes = 0, ws_spmm_bytes = 0;
    uint32_t *tickets = nullptr;
    int *dev_err = nullptr;

    // p2p
    char *peer_arena[kMaxRanks] = {nullptr};
    bool peer_opened[kMaxRanks] = {false};
    bool peer_inproc[kMaxRanks] = {false};  // the peer's engine lives in this process
    bool capturing = false;
    bool connected = false;
    int barrier_next = 0;

    // events (created once)
    cudaEvent_t ev_ready = nullptr, ev_done = nullptr, ev_join = nullptr;
    // input prefetch (set_inputs async == 2): shadow copies filled on the copy stream while an epoch runs
    float *X_shadow = nullptr;
    int32_t *Y_shadow = nullptr;
    float *X_shadow_alloc = nullptr;      // what cudaFree gets: after a flip X_shadow points into the arena
    int32_t *Y_shadow_alloc = nullptr;
    bool flipped = false;                 // X / Y currently are the buffers allocated as shadows (see swap_in_inputs)
    cudaEvent_t ev_pref = nullptr, ev_swap = nullptr;
    bool pref_x = false, pref_y = false, swap_recorded = false;
    std::vector<cudaEvent_t> ev_xfer, ev_spmm;

    // graph
    cudaGraphExec_t graphs[2] = {nullptr, nullptr};  // the captured epoch, per input-buffer assignment (flipped or not)
    float g_lr = 0, g_b1 = 0, g_b2 = 0, g_wd = 0, g_eps = 0;
    int g_skip = 0, g_xs = 0, g_xp = 0;
    int64_t launches_per_epoch = 0;
    bool forward_done = false;

    std::vector<named_timer> timers;
    size_t timer_cursor = 0;
};

namespace {

typedef mggcn_engine eng;

int set_dev(const eng *e) {
    MGGCN_CUDA_TRY(cudaSetDevice(e->device));
    return MGGCN_OK;
}

template <typename T>
T *carve(char *&cur, size_t count) {
    T *p = reinterpret_cast<T *>(cur);
    cur += up256(count * sizeof(T));
    return p;
}

// timers: in eager mode every named region is bracketed by two events created on first use;
// regions may nest and may sit on different streams, so begin hands out a slot and end takes it
int timer_begin(eng *e, const std::string &name, cudaStream_t st, int *slot) {
    *slot = -1;
    if (!e->timers_on) return MGGCN_OK;
    if (e->timer_cursor == e->timers.size()) {
        named_timer t;
        t.name = name;
        MGGCN_CUDA_TRY(cudaEventCreate(&t.beg));
        MGGCN_CUDA_TRY(cudaEventCreate(&t.end));
        e->timers.push_back(t);
    }
    *slot = (int)e->timer_cursor++;
    MGGCN_CUDA_TRY(cudaEventRecord(e->timers[*slot].beg, st));
    return MGGCN_OK;
}
int timer_end(eng *e, cudaStream_t st, int slot, bool spmm = false) {
    if (slot < 0) return MGGCN_OK;
    MGGCN_CUDA_TRY(cudaEventRecord(e->timers[slot].end, st));
    if (spmm) e->timers[slot].cfg = mggcn_last_spmm_config();
    return MGGCN_OK;
}

int xbarrier(eng *e, cudaStream_t st) {
    const int slot = e->barrier_next++;
    if (slot >= kBarrierSlots) return fail(MGGCN_E_STATE, "too many barriers in one epoch");
    peer_table pt{};
    for (int r = 0; r < e->P; ++r) pt.flags[r] = reinterpret_cast<uint32_t *>(e->peer_arena[r]);
    // about 10 s by default; MGGCN_BARRIER_TIMEOUT_MS shortens it for tests.  clock64() ticks at the SM clock.
    static const long long cycles = [] {
        const char *v = getenv("MGGCN_BARRIER_TIMEOUT_MS");
        const long long ms = v && atoll(v) > 0 ? atoll(v) : 10000;
        int dev = 0, khz = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev) != cudaSuccess || khz <= 0) {
            cudaGetLastError();
            khz = 2000000;
        }
        return ms * (long long)khz;
    }();
    xbarrier_kernel<<<1, 32, 0, st>>>(pt, e->tickets, slot, e->rank, e->P, e->dev_err, cycles);
    MGGCN_LAUNCH_CHECK();
    return MGGCN_OK;
}

// One block of a peer's B into `dst`, ordered on the exchange stream e->sm.  The block is cut into g_xstreams
// row ranges that travel on separate streams (a copy-engine transfer is one engine's work however large it is;
// two engines pulling halves of the block roughly double the NVLink rate a rank sees): fork from e->sm, join
// back into it, so the caller's events and barriers on e->sm cover every part.
int peer_pull(eng *e, int blk, size_t off, float *dst, size_t rows, size_t ld) {
    if (g_skip_exchange) return MGGCN_OK;  // measurement only
    const char *src = e->peer_arena[blk] + off;
    const bool by_sm = g_xpull || (e->peer_inproc[blk] && !e->capturing);
    // Large blocks travel as two halves on two engines (one engine does not fill the link: products- and
    // papers100M-shape, 2 GPUs at reddit shape); small blocks of a many-rank exchange, where several pulls are in
    // flight anyway, are faster whole (8 GPUs, reddit shape, 15 MB blocks: 3.18 ms/epoch with 1 stream, 3.31 with 2,
    // 3.82 with 4; profiles/r02_bench_reddit_n8_e2e_sweep.json)
    const int want = g_xstreams > 0 ? g_xstreams : (rows * ld * sizeof(float) <= ((size_t)16 << 20) ? 1 : 2);
    int nx = std::min(want, (int)mggcn_engine::kXStreams);
    nx = std::max(1, std::min(nx, (int)(rows / 1024) + 1));
    if (e->sm == e->sc || by_sm) nx = 1;  // no overlap: everything on the compute stream; SM pulls fill the link alone
    if (nx > 1) MGGCN_CUDA_TRY(cudaEventRecord(e->ev_xfork, e->sm));
    for (int i = nx - 1; i >= 0; --i) {
        const size_t r_lo = rows * (size_t)i / nx, r_hi = rows * (size_t)(i + 1) / nx;
        const size_t count = (r_hi - r_lo) * ld;
        if (!count) continue;
        cudaStream_t st = i == 0 ? e->sm : e->sx[i];
        if (i > 0) MGGCN_CUDA_TRY(cudaStreamWaitEvent(st, e->ev_xfork, 0));
        const float *s_ = reinterpret_cast<const float *>(src) + r_lo * ld;
        float *d_ = dst + r_lo * ld;
        if (by_sm) {
            // count is a multiple of 4 floats (ld is), both ends are 16-byte aligned (256-byte arena carvings, ld % 4 == 0)
            pull_copy_kernel<<<64, 512, 0, st>>>(reinterpret_cast<const float4 *>(s_), reinterpret_cast<float4 *>(d_), count / 4);
            MGGCN_LAUNCH_CHECK();
        } else {
            MGGCN_CUDA_TRY(cudaMemcpyAsync(d_, s_, count * sizeof(float), cudaMemcpyDeviceToDevice, st));
        }
        if (i > 0) {
            MGGCN_CUDA_TRY(cudaEventRecord(e->ev_xjoin[i], st));
            MGGCN_CUDA_TRY(cudaStreamWaitEvent(e->sm, e->ev_xjoin[i], 0));
        }
    }
    return MGGCN_OK;
}

// C = A_which[rank][:] * B  (+ epilogue), B and C are this rank's row blocks (n x ld).
// P stages with a double-buffered exchange; src/cuda_utils.hpp:57-92.
int dist_spmm(eng *e, int which, const float *B, float *C, int64_t F, const mggcn_epilogue *epi,
              const std::string &tname) {
    const int64_t ld = up4(F);
    const int P = e->P;
    if (P == 1) {
        const adj_block &a = e->adj[which][0];
        int ts;
        TRY(timer_begin(e, tname + "matmul-spmm", e->sc, &ts));
        TRY(mggcn_spmm_csr_f32_i32(e->sc, a.rows, a.cols, a.rowptr, a.colidx, a.vals, B, ld, C, ld, (int)F, 1.f, 0.f,
                                   epi, a.plan, e->ws_spmm));
        TRY(timer_end(e, e->sc, ts, true));
        return MGGCN_OK;
    }
    const size_t off = (size_t)(reinterpret_cast<const char *>(B) - e->arena);
    if (e->grouped) {
        // exchange stream: barrier (every rank's B is complete), then pull block (rank+s)%P into its place of the
        // all-gather buffer, s = 1..P-1, with an event behind the last block of every group.  Compute stream: the
        // own block at once, then one plain launch per group of stages (global columns into the all-gather
        // buffer) behind that group's event.
        int t_all, t_x, t_s;
        MGGCN_CUDA_TRY(cudaEventRecord(e->ev_ready, e->sc));
        if (e->sm != e->sc) MGGCN_CUDA_TRY(cudaStreamWaitEvent(e->sm, e->ev_ready, 0));
        TRY(xbarrier(e, e->sm));
        TRY(timer_begin(e, tname + "matmul-spmm", e->sc, &t_all));
        const int G = (int)e->groups.size();
        for (int g = 0; g < G; ++g) {
            const adj_block &a = e->adj[which][g];
            for (int s = std::max(1, e->groups[g].first); s < e->groups[g].second; ++s) {
                const int blk = (e->rank + s) % P;
                float *dst = e->gather + (size_t)e->cuts[blk] * (size_t)ld;
                TRY(timer_begin(e, tname + std::to_string(s) + "_matmul-bcast", e->sm, &t_x));
                TRY(peer_pull(e, blk, off, dst, (size_t)e->cuts[blk + 1] - (size_t)e->cuts[blk], (size_t)ld));
                TRY(timer_end(e, e->sm, t_x));
            }
            if (g > 0 && e->sm != e->sc) {
                MGGCN_CUDA_TRY(cudaEventRecord(e->ev_xfer[g], e->sm));
                MGGCN_CUDA_TRY(cudaStreamWaitEvent(e->sc, e->ev_xfer[g], 0));
            }
            TRY(timer_begin(e, tname + std::to_string(g) + "_matmul-spmm-stage", e->sc, &t_s));
            int64_t touched = 0;  // a group's CSR addresses the whole all-gather buffer but reads its own blocks only
            for (int s = e->groups[g].first; s < e->groups[g].second; ++s)
                touched += (int64_t)e->cuts[(e->rank + s) % P + 1] - (int64_t)e->cuts[(e->rank + s) % P];
            spmm_touched_cols_hint(touched);
            const int rc_g = mggcn_spmm_csr_f32_i32(e->sc, a.rows, a.cols, a.rowptr, a.colidx, a.vals, g == 0 ? B : e->gather, ld,
                                                    C, ld, (int)F, 1.f, g == 0 ? 0.f : 1.f, g == G - 1 ? epi : nullptr, a.plan,
                                                    e->ws_spmm);
            spmm_touched_cols_hint(-1);
            TRY(rc_g);
            TRY(timer_end(e, e->sc, t_s, true));
        }
        TRY(timer_end(e, e->sc, t_all));
        // nobody may overwrite B before every peer has pulled it
        TRY(xbarrier(e, e->sm));
        if (e->sm != e->sc) {
            MGGCN_CUDA_TRY(cudaEventRecord(e->ev_done, e->sm));
            MGGCN_CUDA_TRY(cudaStreamWaitEvent(e->sc, e->ev_done, 0));
        }
        return MGGCN_OK;
    }
    if (e->fused) {
        // compute stream: clear the stage flags, multiply the own block (no peer data needed) -- meanwhile the
        // exchange stream passes the barrier (every rank's B is complete), pulls block (rank+s)%P into its place of
        // the all-gather buffer and publishes flag s behind it.  The remote part of the block row is ONE launch,
        // issued once the barrier has passed: it waits on the device for a stage only if its block has not landed.
        const adj_block &a = e->adj[which][0];
        const adj_block &own = e->adj[which][1];
        int t_all, t_x;
        MGGCN_CUDA_TRY(cudaMemsetAsync(e->stage_flags, 0, 16 * sizeof(uint32_t), e->sc));
        MGGCN_CUDA_TRY(cudaEventRecord(e->ev_ready, e->sc));
        if (e->sm != e->sc) MGGCN_CUDA_TRY(cudaStreamWaitEvent(e->sm, e->ev_ready, 0));
        TRY(xbarrier(e, e->sm));
        if (e->sm != e->sc) MGGCN_CUDA_TRY(cudaEventRecord(e->ev_bar, e->sm));
        TRY(timer_begin(e, tname + "matmul-spmm", e->sc, &t_all));
        TRY(mggcn_spmm_csr_f32_i32(e->sc, own.rows, own.cols, own.rowptr, own.colidx, own.vals, B, ld, C, ld, (int)F, 1.f,
                                   0.f, nullptr, own.plan, e->ws_spmm));
        // the barrier kernel must not queue behind a grid of waiting CTAs: the gathered launch starts after it
        if (e->sm != e->sc) MGGCN_CUDA_TRY(cudaStreamWaitEvent(e->sc, e->ev_bar, 0));
        bool sm_copies = false;
        for (int s = 1; s < P; ++s) {
            const int blk = (e->rank + s) % P;
            const size_t rows_b = (size_t)e->cuts[blk + 1] - (size_t)e->cuts[blk];
            const size_t count = rows_b * (size_t)ld;
            float *dst = e->gather + (size_t)e->cuts[blk] * (size_t)ld;
            TRY(timer_begin(e, tname + std::to_string(s) + "_matmul-bcast", e->sm, &t_x));
            if (e->peer_inproc[blk] && !e->capturing) {
                pull_copy_kernel<<<64, 512, 0, e->sm>>>(reinterpret_cast<const float4 *>(e->peer_arena[blk] + off),
                                                        reinterpret_cast<float4 *>(dst), count / 4);
                MGGCN_LAUNCH_CHECK();
                sm_copies = true;
            } else {
                MGGCN_CUDA_TRY(cudaMemcpyAsync(dst, e->peer_arena[blk] + off, count * sizeof(float),
                                               cudaMemcpyDeviceToDevice, e->sm));
            }
            TRY(publish_flag(e->sm, e->stage_flags + s, e->capturing));
            TRY(timer_end(e, e->sm, t_x));
        }
        if (sm_copies || e->sm == e->sc) {
            // copies done by SMs (eager mode with in-process peers) cannot share the GPU with CTAs that wait
            // for them: finish the transfers first
            if (e->sm != e->sc) {
                MGGCN_CUDA_TRY(cudaEventRecord(e->ev_copied, e->sm));
                MGGCN_CUDA_TRY(cudaStreamWaitEvent(e->sc, e->ev_copied, 0));
            }
        }
        spmm_gather g;
        g.flags = e->stage_flags;
        g.B_own = nullptr;  // the own block has been multiplied already: the rows hold remote columns only
        g.own_lo = e->cuts[e->rank];
        g.own_n = e->cuts[e->rank + 1] - e->cuts[e->rank];
        g.P = P;
        g.rank = e->rank;
        g.err = e->dev_err;
        for (int i = 0; i < 16; ++i) g.cuts[i] = i < P ? e->cuts[i] : 0xFFFFFFFFu;
        TRY(spmm_gathered(e->sc, a.rows, a.cols, a.rowptr, a.colidx, a.vals, e->gather, ld, C, ld, (int)F, 1.f, 1.f, epi,
                          a.plan, e->ws_spmm, g));
        TRY(timer_end(e, e->sc, t_all));
        // nobody may overwrite B before every peer has pulled it
        TRY(xbarrier(e, e->sm));
        if (e->sm != e->sc) {
            MGGCN_CUDA_TRY(cudaEventRecord(e->ev_done, e->sm));
            MGGCN_CUDA_TRY(cudaStreamWaitEvent(e->sc, e->ev_done, 0));
        }
        return MGGCN_OK;
    }
    MGGCN_CUDA_TRY(cudaEventRecord(e->ev_ready, e->sc));
    if (e->sm != e->sc) MGGCN_CUDA_TRY(cudaStreamWaitEvent(e->sm, e->ev_ready, 0));
    TRACE(e->rank, "dist_spmm: events recorded");
    if (e->p2p) TRY(xbarrier(e, e->sm));  // every rank's B is complete
    TRACE(e->rank, "dist_spmm: barrier launched");
    int t_all, t_x, t_s;
    TRY(timer_begin(e, tname + "matmul-spmm", e->sc, &t_all));
    for (int s = 0; s < P; ++s) {
        const int blk = e->p2p ? (e->rank + s) % P : s;
        const adj_block &a = e->adj[which][blk];
        const float *src = B;
        if (blk != e->rank || !e->p2p) {
            // WAR: staging buffer s%2 was last read by the SpMM of stage s-2
            if (s >= 2 && e->sm != e->sc) MGGCN_CUDA_TRY(cudaStreamWaitEvent(e->sm, e->ev_spmm[s - 2], 0));
            const size_t count = (size_t)a.cols * (size_t)ld;
            TRY(timer_begin(e, tname + std::to_string(s) + "_matmul-bcast", e->sm, &t_x));
            if (e->p2p) {
                TRY(peer_pull(e, blk, off, e->stage[s & 1], (size_t)a.cols, (size_t)ld));
                src = e->stage[s & 1];
                TRACE(e->rank, "dist_spmm: peer copy enqueued");
            } else if (blk == e->rank) {
                NCCL_TRY(g_nccl.Broadcast(B, const_cast<float *>(B), count, ncclFloat, blk, e->comm, e->sm));
            } else {
                NCCL_TRY(g_nccl.Broadcast(e->stage[s & 1], e->stage[s & 1], count, ncclFloat, blk, e->comm, e->sm));
                src = e->stage[s & 1];
            }
            TRY(timer_end(e, e->sm, t_x));
            if (e->sm != e->sc) {
                MGGCN_CUDA_TRY(cudaEventRecord(e->ev_xfer[s], e->sm));
                MGGCN_CUDA_TRY(cudaStreamWaitEvent(e->sc, e->ev_xfer[s], 0));
            }
        }
        const bool last = s == P - 1;
        TRY(timer_begin(e, tname + std::to_string(s) + "_matmul-spmm-stage", e->sc, &t_s));
        TRY(mggcn_spmm_csr_f32_i32(e->sc, a.rows, a.cols, a.rowptr, a.colidx, a.vals, src, ld, C, ld, (int)F, 1.f,
                                   s == 0 ? 0.f : 1.f, last ? epi : nullptr, a.plan, e->ws_spmm));
        TRY(timer_end(e, e->sc, t_s, true));
        TRACE(e->rank, "dist_spmm: stage spmm launched");
        if (e->sm != e->sc) MGGCN_CUDA_TRY(cudaEventRecord(e->ev_spmm[s], e->sc));
    }
    TRY(timer_end(e, e->sc, t_all));
    if (e->p2p) {
        // nobody may overwrite B before every peer has pulled it
        TRY(xbarrier(e, e->sm));
        if (e->sm != e->sc) {
            MGGCN_CUDA_TRY(cudaEventRecord(e->ev_done, e->sm));
            MGGCN_CUDA_TRY(cudaStreamWaitEvent(e->sc, e->ev_done, 0));
        }
    }
    return MGGCN_OK;
}

int gemm(eng *e, int ta, int tb, int64_t M, int64_t N, int64_t K, const float *A, int64_t lda, const float *B,
         int64_t ldb, float *C, int64_t ldc, const mggcn_epilogue *epi) {
    return mggcn_gemm_f32(e->sc, ta, tb, M, N, K, 1.f, A, lda, B, ldb, 0.f, C, ldc, epi, e->ws_gemm, e->ws_gemm_bytes);
}

int forward(eng *e) {
    int tg;
    TRACE(e->rank, "forward: begin");
    const float *H = e->X;
    int64_t ldH = e->ldX;
    for (int l = 0; l < e->L; ++l) {
        layer_t &y = e->layers[l];
        const std::string nm = std::to_string(l) + "_0_";
        if (y.Wpad) {
            const int64_t total = y.in * y.ldw;
            pad_rows_kernel<<<(unsigned)std::min<int64_t>((total + 255) / 256, 148 * 4), 256, 0, e->sc>>>(y.W, y.in, y.out, y.Wpad, y.ldw);
            MGGCN_LAUNCH_CHECK();
        }
        mggcn_epilogue act{};
        act.kind = y.activation ? MGGCN_EPI_LEAKY_RELU : MGGCN_EPI_NONE;
        act.slope = e->slope;
        if (y.gemm_first) {
            // HW = H W + b (bias in the GEMM epilogue); Z = act(A HW)
            mggcn_epilogue bias{};
            bias.bias = y.b;
            TRY(timer_begin(e, nm + "matmul-gemm", e->sc, &tg));
            TRY(gemm(e, 0, 0, e->n, y.out, y.in, H, ldH, y.Wt, y.ldw, e->HW, up4(y.out), &bias));
            TRY(timer_end(e, e->sc, tg));
            TRY(dist_spmm(e, 0, e->HW, y.Z, y.out, y.activation ? &act : nullptr, nm));
        } else {
            // AH = A H; Z = act(AH W + b)
            float *AH = y.AH ? y.AH : e->HW;
            TRY(dist_spmm(e, 0, H, AH, y.in, nullptr, nm));
            act.bias = y.b;
            TRY(timer_begin(e, nm + "matmul-gemm", e->sc, &tg));
            TRY(gemm(e, 0, 0, e->n, y.out, y.in, AH, up4(y.in), y.Wt, y.ldw, y.Z, up4(y.out), &act));
            TRY(timer_end(e, e->sc, tg));
        }
        H = y.Z;
        ldH = up4(y.out);
    }
    layer_t &last = e->layers[e->L - 1];
    TRY(timer_begin(e, std::to_string(e->L) + "_loss-layer", e->sc, &tg));
    TRY(mggcn_softmax_xent_f32(e->sc, last.Z, up4(last.out), e->n, (int)last.out, e->Y, 1.f / (float)e->N,
                               e->loss_out, e->ws_loss, e->ws_loss_bytes));
    TRY(timer_end(e, e->sc, tg));
    TRACE(e->rank, "forward: end");
    return MGGCN_OK;
}

int backward(eng *e) {
    int tg;
    const float *T = e->layers[e->L - 1].Z;  // (softmax - onehot)/N, written in place by the loss
    for (int l = e->L - 1; l >= 0; --l) {
        layer_t &y = e->layers[l];
        const bool first = l == 0;
        const std::string nm = std::to_string(l) + "_1_";
        const float *Hin = first ? e->X : e->layers[l - 1].Z;
        const int64_t ldHin = first ? e->ldX : up4(e->layers[l - 1].out);
        float *Tprev = first ? nullptr : e->G[l & 1];
        mggcn_epilogue dact{};  // gradient of layer l-1's LeakyReLU, fused into the producer of T_{l-1}
        if (!first) {
            dact.kind = MGGCN_EPI_LEAKY_RELU_BWD;
            dact.slope = e->slope;
            dact.aux = e->layers[l - 1].Z;
            dact.ld_aux = up4(e->layers[l - 1].out);
        }
        const int64_t ldo = up4(y.out), ldi = up4(y.in);
        if (y.gemm_first) {
            const float *GHW = T;
            if (!first || e->exact) {  // the reference skips this SpMM for layer 0
                TRY(dist_spmm(e, 1, T, e->HW, y.out, nullptr, nm));
                GHW = e->HW;
            }
            TRY(timer_begin(e, nm + "matmul-gemm", e->sc, &tg));
            TRY(mggcn_colsum_f32(e->sc, e->n, y.out, GHW, ldo, y.Gb, e->ws_colsum, e->ws_colsum_bytes));
            TRY(gemm(e, 1, 0, y.in, y.out, e->n, Hin, ldHin, GHW, ldo, y.GW, y.out, nullptr));
            if (!first) TRY(gemm(e, 0, 1, e->n, y.in, y.out, GHW, ldo, y.Wt, y.ldw, Tprev, ldi, &dact));
            TRY(timer_end(e, e->sc, tg));
        } else {
            // the reference uses the raw layer input H as the GEMM input here
            const float *Xin = e->exact ? y.AH : Hin;
            const int64_t ldXin = e->exact ? ldi : ldHin;
            TRY(timer_begin(e, nm + "matmul-gemm", e->sc, &tg));
            TRY(mggcn_colsum_f32(e->sc, e->n, y.out, T, ldo, y.Gb, e->ws_colsum, e->ws_colsum_bytes));
            TRY(gemm(e, 1, 0, y.in, y.out, e->n, Xin, ldXin, T, ldo, y.GW, y.out, nullptr));
            if (!first) TRY(gemm(e, 0, 1, e->n, y.in, y.out, T, ldo, y.Wt, y.ldw, e->HW, ldi, nullptr));
            TRY(timer_end(e, e->sc, tg));
            if (!first) TRY(dist_spmm(e, 1, e->HW, Tprev, y.in, &dact, nm));
        }
        T = Tprev;
    }
    // weight-gradient reduction over ranks (reference: one ncclAllReduce per tensor,
    // src/dist_matrix.hpp:587-592 via src/gcn.hpp:238-240)
    if (e->P > 1) {
        TRY(timer_begin(e, "allreduce", e->sc, &tg));
        if (e->p2p) {
            MGGCN_CUDA_TRY(cudaEventRecord(e->ev_ready, e->sc));
            if (e->sm != e->sc) MGGCN_CUDA_TRY(cudaStreamWaitEvent(e->sm, e->ev_ready, 0));
            TRY(xbarrier(e, e->sm));
            if (e->sm != e->sc) {
                MGGCN_CUDA_TRY(cudaEventRecord(e->ev_done, e->sm));
                MGGCN_CUDA_TRY(cudaStreamWaitEvent(e->sc, e->ev_done, 0));
            }
            grad_peers gp{};
            const size_t off = (size_t)(reinterpret_cast<char *>(e->grads) - e->arena);
            for (int r = 0; r < e->P; ++r) gp.g[r] = reinterpret_cast<const float *>(e->peer_arena[r] + off);
            grad_reduce_kernel<<<(unsigned)std::max<size_t>(1, std::min<size_t>((e->n_params + 255) / 256, 296)), 256, 0,
                                 e->sc>>>(gp, e->P, e->n_params, e->gsum);
            MGGCN_LAUNCH_CHECK();
        } else {
            NCCL_TRY(g_nccl.AllReduce(e->grads, e->gsum, e->n_params, ncclFloat, ncclSum, e->comm, e->sc));
        }
        TRY(timer_end(e, e->sc, tg));
    }
    return MGGCN_OK;
}

int adam(eng *e, float lr, float b1, float b2, float wd, float eps) {
    int tg;
    TRY(timer_begin(e, "adam-update", e->sc, &tg));
    adam_prep_kernel<<<1, 1, 0, e->sc>>>(e->adam_state, b1, b2);
    MGGCN_LAUNCH_CHECK();
    const unsigned grid = (unsigned)std::max<size_t>(1, std::min<size_t>((e->n_params + 255) / 256, 148 * 4));
    adam_dev_kernel<<<grid, 256, 0, e->sc>>>(e->n_params, e->params, e->gsum, e->adam_m, e->adam_v, e->adam_state, lr,
                                             b1, b2, wd, eps, e->n_weights, e->dev_err);
    MGGCN_LAUNCH_CHECK();
    TRY(timer_end(e, e->sc, tg));
    return MGGCN_OK;
}

void drop_graphs(eng *e) {
    for (cudaGraphExec_t &g : e->graphs)
        if (g) { cudaGraphExecDestroy(g); g = nullptr; }
}

// Makes prefetched inputs current: ordered after the prefetch on the copy stream and after everything
// already queued on the compute stream (the previous epoch's backward reads X).
int swap_in_inputs(eng *e) {
    if (!e->pref_x && !e->pref_y) return MGGCN_OK;
    MGGCN_CUDA_TRY(cudaStreamWaitEvent(e->sc, e->ev_pref, 0));
    if (g_input_flip && e->P == 1 && e->pref_x && e->pref_y) {
        // single-GPU engine, both inputs prefetched: the shadows BECOME the inputs (no peer addresses X by arena offset
        // here), and the epoch graph exists once per assignment -- saves a device-to-device copy of X per step
        // (567 MB read + written = 0.4 ms of a 20 ms epoch at reddit shape).  The next prefetch lands in the old
        // buffers, behind ev_swap, i.e. after everything queued so far has stopped reading them.
        std::swap(e->X, e->X_shadow);
        std::swap(e->Y, e->Y_shadow);
        e->flipped = !e->flipped;
    } else {
        if (e->pref_x)
            MGGCN_CUDA_TRY(cudaMemcpyAsync(e->X, e->X_shadow, (size_t)e->n * (size_t)e->ldX * 4, cudaMemcpyDeviceToDevice, e->sc));
        if (e->pref_y) MGGCN_CUDA_TRY(cudaMemcpyAsync(e->Y, e->Y_shadow, (size_t)e->n * 4, cudaMemcpyDeviceToDevice, e->sc));
    }
    MGGCN_CUDA_TRY(cudaEventRecord(e->ev_swap, e->sc));
    e->swap_recorded = true;
    e->pref_x = e->pref_y = false;
    return MGGCN_OK;
}

int check_ready(eng *e) {
    if (!e->adj_set[0] || !e->adj_set[1]) return fail(MGGCN_E_STATE, "adjacency not set");
    if (e->P > 1 && e->p2p && !e->connected) return fail(MGGCN_E_STATE, "p2p engine not connected to its peers");
    return MGGCN_OK;
}

void begin_pass(eng *e) {
    e->barrier_next = 0;
    e->timer_cursor = 0;
}

int epoch_body(eng *e, float lr, float b1, float b2, float wd, float eps) {
    begin_pass(e);
    TRY(forward(e));
    TRY(backward(e));
    TRY(adam(e, lr, b1, b2, wd, eps));
    return MGGCN_OK;
}

}  // namespace

extern "C" {

int mggcn_nccl_unique_id(void *id128) {
    if (!id128) return fail(MGGCN_E_INVALID, "null id");
    TRY(nccl_load());
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    NCCL_TRY(g_nccl.GetUniqueId(static_cast<ncclUniqueId *>(id128)));
    return MGGCN_OK;
}

int mggcn_engine_create(const mggcn_engine_config *cfg, mggcn_engine_t **out) {
    if (!cfg || !out) return fail(MGGCN_E_INVALID, "null argument");
    if (cfg->world < 1 || cfg->world > kMaxRanks || cfg->rank < 0 || cfg->rank >= cfg->world)
        return fail(MGGCN_E_INVALID, "bad world/rank %d/%d", cfg->world, cfg->rank);
    if (cfg->num_layers < 1 || !cfg->sizes || !cfg->p) return fail(MGGCN_E_INVALID, "bad layer description");
    auto *e = new (std::nothrow) mggcn_engine;
    if (!e) return fail(MGGCN_E_NOMEM, "out of memory");
    e->P = cfg->world;
    e->rank = cfg->rank;
    e->device = cfg->device;
    e->L = cfg->num_layers;
    e->overlap = cfg->overlap != 0;
    e->exact = cfg->exact_grads != 0;
    e->use_graph = cfg->use_graph != 0;
    e->p2p = (cfg->comm_backend == MGGCN_COMM_P2P || cfg->comm_backend == MGGCN_COMM_P2P_STAGED ||
              cfg->comm_backend == MGGCN_COMM_P2P_GROUPED) && cfg->world > 1;
    e->fused = e->p2p && cfg->comm_backend == MGGCN_COMM_P2P && !getenv("MGGCN_STAGED");
    e->grouped = e->p2p && cfg->comm_backend == MGGCN_COMM_P2P_GROUPED;
    // the all-gather buffer holds every rank's rows at the widest layer: past 32 GB per rank (papers100M-shape:
    // 57 GB) the grouped form falls back to the double-buffered staged exchange
    if (e->grouped) {
        int64_t w = 0;  // widest aggregated operand: a layer aggregates its narrower side (src/gcn.hpp:439,558)
        for (int l = 0; l < cfg->num_layers; ++l) w = std::max<int64_t>(w, up4(std::min(cfg->sizes[l], cfg->sizes[l + 1])));
        if ((double)cfg->n_global * (double)w * 4.0 > 32.0 * (1ull << 30)) e->grouped = false;
    }
    if (e->grouped) {  // provisional; choose_groups() decides once the block row is known
        e->groups.push_back({0, 1});
        for (int s0 = 1; s0 < e->P; s0 *= 2) e->groups.push_back({s0, std::min(2 * s0, e->P)});
    }
    e->timers_on = cfg->timers != 0 && !e->use_graph;
    e->slope = cfg->lrelu_slope;
    e->N = cfg->n_global;
    e->cuts.assign(cfg->p, cfg->p + e->P + 1);
    e->n = (int64_t)e->cuts[e->rank + 1] - (int64_t)e->cuts[e->rank];
    e->sizes.assign(cfg->sizes, cfg->sizes + e->L + 1);
    for (int64_t s : e->sizes)
        if (s <= 0) { delete e; return fail(MGGCN_E_INVALID, "layer width must be positive"); }
#define E_TRY(expr)                                  \
    do {                                             \
        int rc__ = (expr);                           \
        if (rc__) { mggcn_engine_destroy(e); return rc__; } \
    } while (0)
#define E_CUDA(expr)                                                                                 \
    do {                                                                                             \
        cudaError_t ce__ = (expr);                                                                   \
        if (ce__ != cudaSuccess) {                                                                   \
            mggcn_engine_destroy(e);                                                                 \
            return fail((int)ce__, "%s: %s", #expr, cudaGetErrorString(ce__));                      \
        }                                                                                            \
    } while (0)
    E_CUDA(cudaSetDevice(e->device));
    int lo = 0, hi = 0;
    E_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    // compute = low priority, exchange = high priority (src/matrix.hpp:53-60,82)
    E_CUDA(cudaStreamCreateWithPriority(&e->sc, cudaStreamNonBlocking, lo));
    if (e->overlap && e->P > 1) E_CUDA(cudaStreamCreateWithPriority(&e->sm, cudaStreamNonBlocking, hi));
    else e->sm = e->sc;
    E_CUDA(cudaStreamCreateWithFlags(&e->scopy, cudaStreamNonBlocking));
    e->sx[0] = e->sm;
    if (e->sm != e->sc) {
        E_CUDA(cudaEventCreateWithFlags(&e->ev_xfork, cudaEventDisableTiming));
        for (int i = 1; i < mggcn_engine::kXStreams; ++i) {
            E_CUDA(cudaStreamCreateWithPriority(&e->sx[i], cudaStreamNonBlocking, hi));
            E_CUDA(cudaEventCreateWithFlags(&e->ev_xjoin[i], cudaEventDisableTiming));
        }
    }
    E_CUDA(cudaEventCreateWithFlags(&e->ev_ready, cudaEventDisableTiming));
    E_CUDA(cudaEventCreateWithFlags(&e->ev_done, cudaEventDisableTiming));
    E_CUDA(cudaEventCreateWithFlags(&e->ev_join, cudaEventDisableTiming));
    E_CUDA(cudaEventCreateWithFlags(&e->ev_bar, cudaEventDisableTiming));
    E_CUDA(cudaEventCreateWithFlags(&e->ev_copied, cudaEventDisableTiming));
    e->ev_xfer.resize(e->P);
    e->ev_spmm.resize(e->P);
    for (int i = 0; i < e->P; ++i) {
        E_CUDA(cudaEventCreateWithFlags(&e->ev_xfer[i], cudaEventDisableTiming));
        E_CUDA(cudaEventCreateWithFlags(&e->ev_spmm[i], cudaEventDisableTiming));
    }

    // ---- layers and sizes
    e->layers.resize(e->L);
    e->n_weights = 0;
    size_t n_bias = 0;
    // widths: `maxw` = widest aggregated operand (gemm-first layers aggregate their output, the others their
    // input) -- it sizes HW, the staging / all-gather buffers and the SpMM workspace; `gw` = widest gradient
    // handed to the layer below (the input width of layers 1..L-1)
    e->maxw = 4;
    e->gw = 4;
    for (int l = 0; l < e->L; ++l) {
        layer_t &y = e->layers[l];
        y.in = e->sizes[l];
        y.out = e->sizes[l + 1];
        y.activation = l + 1 < e->L;      // src/gcn.hpp:954,1023
        y.gemm_first = y.out <= y.in;     // src/gcn.hpp:439,558
        e->n_weights += (size_t)(y.in * y.out);
        n_bias += (size_t)y.out;
        e->maxw = std::max(e->maxw, up4(y.gemm_first ? y.out : y.in));
        if (l > 0) e->gw = std::max(e->gw, up4(y.in));
    }
    e->n_params = e->n_weights + n_bias;
    for (int i = 0; i < e->P; ++i)
        e->max_block_rows = std::max<int64_t>(e->max_block_rows, (int64_t)e->cuts[i + 1] - (int64_t)e->cuts[i]);
    e->ldX = up4(e->sizes[0]);

    // Every per-row buffer is carved for the LARGEST row block of the partition, not for this rank's own: peers
    // address each other's buffers as (peer arena base + this rank's offset), so the layout must not depend on
    // the rank even when N % P != 0 makes the blocks uneven.
    for (int l = 0; l < e->L; ++l) {
        const layer_t &y = e->layers[l];
        // the three products of a layer share one workspace: forward H*W (pre-split W), dW = H^T*G (split-K partials),
        // dH = G*W^T (pre-split W); a workspace sized for dW alone made the other two fall back to in-kernel splitting
        // whenever the partials happened to be the smaller need
        e->ws_gemm_bytes = std::max({e->ws_gemm_bytes,
                                     mggcn_gemm_workspace_bytes(0, 0, e->max_block_rows, y.out, y.in),
                                     mggcn_gemm_workspace_bytes(1, 0, y.in, y.out, e->max_block_rows),
                                     mggcn_gemm_workspace_bytes(0, 1, e->max_block_rows, y.in, y.out)});
        e->ws_colsum_bytes = std::max(e->ws_colsum_bytes, mggcn_colsum_workspace_bytes(e->max_block_rows, y.out));
    }
    e->ws_loss_bytes = mggcn_softmax_xent_workspace_bytes(e->max_block_rows);

    // ---- arena
    const size_t n = (size_t)e->max_block_rows;
    size_t total = 0;
    auto need = [&](size_t bytes) { total += up256(bytes); };
    need(sizeof(uint32_t) * kBarrierSlots * kMaxRanks);           // flags
    need(n * (size_t)e->ldX * 4);                                  // X
    need(n * 4);                                                   // Y
    need(n * (size_t)e->maxw * 4);                                 // HW
    need(n * (size_t)e->gw * 4);                                   // G0
    need(n * (size_t)e->gw * 4);                                   // G1
    if (e->fused || e->grouped) {
        need((size_t)e->N * (size_t)e->maxw * 4);
        need(64);
    } else if (e->P > 1) {
        need((size_t)e->max_block_rows * (size_t)e->maxw * 4);
        need((size_t)e->max_block_rows * (size_t)e->maxw * 4);
    }
    for (int l = 0; l < e->L; ++l) {
        need(n * (size_t)up4(e->layers[l].out) * 4);
        if (e->exact && !e->layers[l].gemm_first) need(n * (size_t)up4(e->layers[l].in) * 4);
    }
    for (int k = 0; k < 5; ++k) need(e->n_params * 4);  // params grads gsum m v
    need(16);                                           // adam_state
    need(16);                                           // loss_out
    need(e->ws_gemm_bytes);
    need(e->ws_colsum_bytes);
    need(e->ws_loss_bytes);
    need(sizeof(uint32_t) * kBarrierSlots);
    need(16);
    e->arena_bytes = total + 4096;
    E_CUDA(cudaMalloc(&e->arena, e->arena_bytes));
    E_CUDA(cudaMemset(e->arena, 0, e->arena_bytes));
    char *cur = e->arena;
    e->flags = carve<uint32_t>(cur, kBarrierSlots * kMaxRanks);
    e->X = carve<float>(cur, n * (size_t)e->ldX);
    e->Y = carve<int32_t>(cur, n);
    e->HW = carve<float>(cur, n * (size_t)e->maxw);
    e->G[0] = carve<float>(cur, n * (size_t)e->gw);
    e->G[1] = carve<float>(cur, n * (size_t)e->gw);
    if (e->fused || e->grouped) {
        e->gather = carve<float>(cur, (size_t)e->N * (size_t)e->maxw);
        e->stage_flags = carve<uint32_t>(cur, 16);
    } else if (e->P > 1) {
        e->stage[0] = carve<float>(cur, (size_t)e->max_block_rows * (size_t)e->maxw);
        e->stage[1] = carve<float>(cur, (size_t)e->max_block_rows * (size_t)e->maxw);
    }
    for (int l = 0; l < e->L; ++l) {
        e->layers[l].Z = carve<float>(cur, n * (size_t)up4(e->layers[l].out));
        if (e->exact && !e->layers[l].gemm_first) e->layers[l].AH = carve<float>(cur, n * (size_t)up4(e->layers[l].in));
    }
    e->params = carve<float>(cur, e->n_params);
    e->grads = carve<float>(cur, e->n_params);
    e->gsum = e->P > 1 ? carve<float>(cur, e->n_params) : e->grads;
    if (e->P == 1) carve<float>(cur, e->n_params);
    e->adam_m = carve<float>(cur, e->n_params);
    e->adam_v = carve<float>(cur, e->n_params);
    e->adam_state = carve<float>(cur, 4);
    e->loss_out = carve<double>(cur, 2);
    e->ws_gemm = e->ws_gemm_bytes ? carve<char>(cur, e->ws_gemm_bytes) : nullptr;
    e->ws_colsum = carve<char>(cur, e->ws_colsum_bytes);
    e->ws_loss = carve<char>(cur, e->ws_loss_bytes);
    e->tickets = carve<uint32_t>(cur, kBarrierSlots);
    e->dev_err = carve<int>(cur, 4);
    e->peer_arena[e->rank] = e->arena;

    // ---- parameters: packed [W_0 .. W_{L-1} | b_0 .. b_{L-1}], reference init (src/gcn.hpp:211-214)
    {
        std::vector<float> host(e->n_params);
        size_t wo = 0, bo = e->n_weights;
        for (int l = 0; l < e->L; ++l) {
            layer_t &y = e->layers[l];
            y.W = e->params + wo;
            y.GW = e->grads + wo;
            y.b = e->params + bo;
            y.Gb = e->grads + bo;
            E_TRY(mggcn_init_weights(y.in, y.out, (float)std::sqrt(2.0 / (1.0 + 0.01 * 0.01)), host.data() + wo));
            E_TRY(mggcn_init_weights(1, y.out, std::sqrt(1.f / 3.f), host.data() + bo));
            wo += (size_t)(y.in * y.out);
            bo += (size_t)y.out;
        }
        E_CUDA(cudaMemcpy(e->params, host.data(), e->n_params * 4, cudaMemcpyDefault));
        for (int l = 0; l < e->L; ++l) {
            layer_t &y = e->layers[l];
            y.Wt = y.W;
            y.ldw = y.out;
            if (y.out % 4 != 0 || (reinterpret_cast<uintptr_t>(y.W) & 15) != 0) {
                y.ldw = up4(y.out);
                E_CUDA(cudaMalloc(&y.Wpad, (size_t)(y.in * y.ldw) * sizeof(float)));
                y.Wt = y.Wpad;
            }
        }
    }

    if (e->P > 1 && !e->p2p) {
        if (!cfg->nccl_id) { mggcn_engine_destroy(e); return fail(MGGCN_E_INVALID, "nccl_id is required when world > 1"); }
        E_TRY(nccl_load());
        ncclUniqueId id;
        memcpy(&id, cfg->nccl_id, sizeof id);
        ncclResult_t r = g_nccl.CommInitRank(&e->comm, e->P, id, e->rank);
        if (r != ncclSuccess) {
            mggcn_engine_destroy(e);
            return fail(MGGCN_E_NCCL, "ncclCommInitRank: %s", g_nccl.GetErrorString(r));
        }
    }
#undef E_TRY
#undef E_CUDA
    *out = e;
    return MGGCN_OK;
}

// 0 = well formed; 1 = rowptr[0] != 0; 2 = rowptr not monotone; 3 = a column id >= cols
static int validate_csr(int64_t rows, int64_t cols, const uint64_t *rp, const uint32_t *ci) {
    if (rp[0] != 0) return 1;
    int bad = 0;
#pragma omp parallel for schedule(static) reduction(max : bad)
    for (int64_t r = 0; r < rows; ++r) {
        if (rp[r + 1] < rp[r]) { bad = std::max(bad, 2); continue; }
        for (uint64_t k = rp[r]; k < rp[r + 1]; ++k)
            if ((int64_t)ci[k] >= cols) { bad = std::max(bad, 3); break; }
    }
    return bad;
}

// Which stages share a launch in the grouped exchange.  Long groups give the kernel longer rows and grids (what an
// 8 x 8 block lacks), but a group starts only when its LAST block has landed: the final group of the doubling
// scheme holds half of the block row behind the whole exchange.  That is free while a block multiplies slower
// than it travels (reddit-, proteins-shape: 60-75 entries per row and block) and costs up to half an aggregation
// when the exchange is the longer side (products-shape at 8 GPUs: 6 entries per row and block, 313 MB blocks
// that miss the L2: 12.4 of 30.3 ms exposed with the doubling scheme).  Estimate both sides per block:
//   transfer = rows_blk * w * 4 / X,  multiply = nnz_blk * w * 4 / R
// with X = 500 GB/s (what a rank receives while every rank pulls and multiplies) and R = the gather rate the SpMM
// sustains (16 TB/s out of the L2, 6.5 TB/s when a block of B is larger than half the L2), i.e. the exchange is
// the longer side when the entries per row and block are fewer than R / X.
// Scope of the rule: X and R are B200 / NVLink-5 figures measured in this repository (DESIGN.md section 4 and
// profiles/l2_peak.json); on another part set the grouping with mggcn_tune("xgroups", 1 | 2) instead.  The decision is
// taken per rank from that rank's own block row, without an exchange: ranks whose rows sit near the threshold may
// group differently, which changes their launch counts but not the protocol -- every rank passes the same number of
// stage barriers and publishes the same flags whatever its grouping (tests/dist_engine_check.py runs mixed groupings).
static void choose_groups(mggcn_engine *e, const uint64_t *const *blk_rowptr) {
    if (!e->grouped || e->groups_chosen) return;
    e->groups_chosen = true;
    bool singles = g_xgroups == 1;
    if (g_xgroups == 0) {
        double nnz = 0;
        for (int j = 0; j < e->P; ++j) nnz += (double)blk_rowptr[j][e->n];
        const double per_row_and_block = nnz / std::max<double>(1.0, (double)e->n * (double)e->P);
        int dev = 0, l2 = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, dev) != cudaSuccess || l2 <= 0) {
            cudaGetLastError();
            l2 = 126 << 20;
        }
        int64_t blk_rows = 0;
        for (int j = 0; j < e->P; ++j) blk_rows = std::max<int64_t>(blk_rows, (int64_t)e->cuts[j + 1] - (int64_t)e->cuts[j]);
        const double blk_bytes = (double)blk_rows * (double)e->maxw * 4.0;
        const double R = blk_bytes > 0.5 * (double)l2 ? 6.5e12 : 16e12, X = 5e11;
        singles = per_row_and_block < R / X;
    }
    e->groups.clear();
    e->groups.push_back({0, 1});
    if (singles)
        for (int s0 = 1; s0 < e->P; ++s0) e->groups.push_back({s0, s0 + 1});
    else
        for (int s0 = 1; s0 < e->P; s0 *= 2) e->groups.push_back({s0, std::min(2 * s0, e->P)});
}

int mggcn_engine_set_adjacency(mggcn_engine_t *e, int which, const uint64_t *const *blk_rowptr,
                               const uint32_t *const *blk_colidx, const float *const *blk_vals) {
    if (!e || which < 0 || which > 1 || !blk_rowptr || !blk_colidx || !blk_vals) return fail(MGGCN_E_INVALID, "bad argument");
    TRY(set_dev(e));
    // the kernels gather B + colidx * ld unchecked: refuse malformed blocks here, once, on the host
    for (int j = 0; j < e->P; ++j) {
        if (!blk_rowptr[j] || ((!blk_colidx[j] || !blk_vals[j]) && blk_rowptr[j][e->n] > 0))
            return fail(MGGCN_E_INVALID, "block %d has a null array", j);
        const int rc = validate_csr(e->n, (int64_t)e->cuts[j + 1] - (int64_t)e->cuts[j], blk_rowptr[j], blk_colidx[j]);
        if (rc) return fail(MGGCN_E_FORMAT, "adjacency block (%d, %d) is malformed: %s", e->rank, j,
                            rc == 1 ? "row pointers do not start at 0" : rc == 2 ? "row pointers decrease" : "column index out of range");
    }
    choose_groups(e, blk_rowptr);
    size_t ws = e->ws_spmm_bytes;
    // merged CSRs over ranges of stages (global column ids into the all-gather buffer, every row in stage order:
    // blocks rank+s0, rank+s0+1, ... cyclically).  Gathered form: slot 0 = all remote stages [1, P) in one CSR, so
    // one launch can follow the transfers; slot 1 (below) = the own block with block-local columns, multiplied
    // first while the barrier and the first transfer run.  Grouped form: slot g = stages of group g (g >= 1),
    // slot 0 (below) = the own block.
    std::vector<std::pair<int, std::pair<int, int>>> merged;  // (slot, [s0, s1))
    if (e->fused) merged.push_back({0, {1, e->P}});
    if (e->grouped)
        for (size_t g = 1; g < e->groups.size(); ++g) merged.push_back({(int)g, e->groups[g]});
    for (const auto &mg : merged) {
        adj_block &a = e->adj[which][mg.first];
        const int s0 = mg.second.first, s1 = mg.second.second;
        if (a.plan) { mggcn_spmm_plan_destroy(a.plan); a.plan = nullptr; }
        if (a.rowptr) { cudaFree(a.rowptr); a.rowptr = nullptr; }
        if (a.colidx) { cudaFree(a.colidx); a.colidx = nullptr; }
        if (a.vals) { cudaFree(a.vals); a.vals = nullptr; }
        const int P = e->P;
        const int64_t n = e->n;
        std::vector<uint64_t> rp((size_t)n + 1, 0);
        for (int64_t r = 0; r < n; ++r) {
            uint64_t c = 0;
            for (int st = s0; st < s1; ++st) {
                const int j = (e->rank + st) % P;
                c += blk_rowptr[j][r + 1] - blk_rowptr[j][r];
            }
            rp[(size_t)r + 1] = rp[(size_t)r] + c;
        }
        const uint64_t nnz = rp[(size_t)n];
        if (nnz > 0xFFFFFFFFull) return fail(MGGCN_E_UNSUPPORTED, "block row has more than 2^32 entries");
        std::vector<uint32_t> ci((size_t)std::max<uint64_t>(nnz, 1));
        std::vector<float> va((size_t)std::max<uint64_t>(nnz, 1));
#pragma omp parallel for schedule(static)
        for (int64_t r = 0; r < n; ++r) {
            uint64_t o = rp[(size_t)r];
            for (int st = s0; st < s1; ++st) {
                const int j = (e->rank + st) % P;
                const uint32_t base = e->cuts[j];
                for (uint64_t k = blk_rowptr[j][r]; k < blk_rowptr[j][r + 1]; ++k, ++o) {
                    ci[(size_t)o] = blk_colidx[j][k] + base;
                    va[(size_t)o] = blk_vals[j][k];
                }
            }
        }
        std::vector<uint32_t> rp32((size_t)n + 1);
        for (int64_t r = 0; r <= n; ++r) rp32[(size_t)r] = (uint32_t)rp[(size_t)r];
        a.rows = n;
        a.cols = e->N;
        a.nnz = (int64_t)nnz;
        MGGCN_CUDA_TRY(cudaMalloc(&a.rowptr, rp32.size() * 4));
        MGGCN_CUDA_TRY(cudaMemcpy(a.rowptr, rp32.data(), rp32.size() * 4, cudaMemcpyDefault));
        MGGCN_CUDA_TRY(cudaMalloc(&a.colidx, ci.size() * 4));
        MGGCN_CUDA_TRY(cudaMalloc(&a.vals, va.size() * 4));
        MGGCN_CUDA_TRY(cudaMemcpy(a.colidx, ci.data(), (size_t)nnz * 4, cudaMemcpyDefault));
        MGGCN_CUDA_TRY(cudaMemcpy(a.vals, va.data(), (size_t)nnz * 4, cudaMemcpyDefault));
        if (e->grouped) {
            // the group's columns are the blocks rank+s0 .. rank+s1-1 (cyclic) of the all-gather buffer; far beyond
            // the L2 (products-shape) the plan cuts them into column blocks (2-D schedule, csrc/spmm.cu)
            int64_t span = 0;
            for (int st = s0; st < s1; ++st) span += (int64_t)e->cuts[(e->rank + st) % P + 1] - (int64_t)e->cuts[(e->rank + st) % P];
            TRY(mggcn_spmm_plan_create_blocked_i32(a.rows, a.cols, rp32.data(), ci.data(), e->cuts[(e->rank + s0) % P], span,
                                                   (int)e->maxw, 0, &a.plan));
        } else {
            TRY(mggcn_spmm_plan_create_i32(a.rows, rp32.data(), 0, &a.plan));
        }
        ws = std::max(ws, mggcn_spmm_plan_workspace_bytes(a.plan, (int)e->maxw));
    }
    for (int j = 0; j < e->P; ++j) {
        if ((e->fused || e->grouped) && j != e->rank) continue;
        adj_block &a = e->adj[which][e->fused ? 1 : e->grouped ? 0 : j];
        if (a.plan) { mggcn_spmm_plan_destroy(a.plan); a.plan = nullptr; }
        if (a.rowptr) { cudaFree(a.rowptr); a.rowptr = nullptr; }
        if (a.colidx) { cudaFree(a.colidx); a.colidx = nullptr; }
        if (a.vals) { cudaFree(a.vals); a.vals = nullptr; }
        a.rows = e->n;
        a.cols = (int64_t)e->cuts[j + 1] - (int64_t)e->cuts[j];
        const uint64_t *rp = blk_rowptr[j];
        a.nnz = (int64_t)rp[a.rows];
        if ((uint64_t)a.nnz > 0xFFFFFFFFull) return fail(MGGCN_E_UNSUPPORTED, "block nnz exceeds 32 bits");
        // after the P x P split every block's local row pointers fit 32 bits
        std::vector<uint32_t> rp32((size_t)a.rows + 1);
        for (int64_t r = 0; r <= a.rows; ++r) rp32[(size_t)r] = (uint32_t)rp[r];
        MGGCN_CUDA_TRY(cudaMalloc(&a.rowptr, rp32.size() * 4));
        MGGCN_CUDA_TRY(cudaMemcpy(a.rowptr, rp32.data(), rp32.size() * 4, cudaMemcpyDefault));
        MGGCN_CUDA_TRY(cudaMalloc(&a.colidx, std::max<size_t>((size_t)a.nnz, 1) * 4));
        MGGCN_CUDA_TRY(cudaMalloc(&a.vals, std::max<size_t>((size_t)a.nnz, 1) * 4));
        MGGCN_CUDA_TRY(cudaMemcpy(a.colidx, blk_colidx[j], (size_t)a.nnz * 4, cudaMemcpyDefault));
        MGGCN_CUDA_TRY(cudaMemcpy(a.vals, blk_vals[j], (size_t)a.nnz * 4, cudaMemcpyDefault));
        if (e->fused) TRY(mggcn_spmm_plan_create_i32(a.rows, rp32.data(), 0, &a.plan));
        else TRY(mggcn_spmm_plan_create_blocked_i32(a.rows, a.cols, rp32.data(), blk_colidx[j], 0, a.cols, (int)e->maxw, 0, &a.plan));
        ws = std::max(ws, mggcn_spmm_plan_workspace_bytes(a.plan, (int)e->maxw));
    }
    if (ws > e->ws_spmm_bytes) {
        if (e->ws_spmm) cudaFree(e->ws_spmm);
        MGGCN_CUDA_TRY(cudaMalloc(&e->ws_spmm, ws));
        e->ws_spmm_bytes = ws;
    }
    e->adj_set[which] = true;
    drop_graphs(e);
    return MGGCN_OK;
}

// host rows (pitch = width floats) into device rows of pitch ld: one flat copy when the pitches agree
// Size of the pieces a contiguous host -> device input transfer is issued in.  On a multi-GPU engine the transfer
// shares the copy engines with the peer pulls of the running epoch; behind ONE 71 MB transfer the pulls of a whole
// epoch queued up (8 GPUs, reddit shape: the epoch ran 7.4 ms instead of 3.3 while a foreign stream copied, the
// pipelined step 5.4 ms).  In 256 KB pieces the engine takes the (higher-priority) pulls in between: 3.8 ms
// (profiles/r02_bench_reddit_n8_e2e_sweep.json).  At most 256 pieces, so the host side stays below a millisecond.
static size_t h2d_piece_bytes(size_t total, int P) {
    if (g_h2d_chunk_kb > 0) return (size_t)g_h2d_chunk_kb << 10;
    if (P <= 1) return total;
    const size_t piece = std::max<size_t>((size_t)256 << 10, (total + 255) / 256);
    return (piece + 65535) & ~(size_t)65535;
}

static int copy_rows_in(float *dst, int64_t ld, const float *src, int64_t width, int64_t rows, cudaStream_t st, int P) {
    if (g_h2d_sm && ld == width && ((size_t)rows * (size_t)width) % 4 == 0 && (reinterpret_cast<uintptr_t>(src) & 15) == 0) {
        // pinned host memory is mapped into the device's address space: a few CTAs reading it with 128-bit loads
        // keep PCIe full and leave the copy engines to the peer exchange (an H2D transfer and the pulls of an
        // aggregation otherwise queue on the same engines, tune key "h2d_sm")
        cudaPointerAttributes attr;
        if (cudaPointerGetAttributes(&attr, src) == cudaSuccess && attr.type == cudaMemoryTypeHost && attr.devicePointer) {
            pull_copy_kernel<<<g_h2d_sm, 512, 0, st>>>(reinterpret_cast<const float4 *>(attr.devicePointer),
                                                     reinterpret_cast<float4 *>(dst), (size_t)rows * (size_t)width / 4);
            MGGCN_LAUNCH_CHECK();
            return MGGCN_OK;
        }
        cudaGetLastError();
    }
    if (ld == width) {
        const size_t total = (size_t)rows * (size_t)width * 4;
        const size_t piece = std::max<size_t>(1, h2d_piece_bytes(total, P));
        for (size_t off = 0; off < total; off += piece)
            MGGCN_CUDA_TRY(cudaMemcpyAsync(reinterpret_cast<char *>(dst) + off, reinterpret_cast<const char *>(src) + off,
                                           std::min(piece, total - off), cudaMemcpyDefault, st));
    } else {
        MGGCN_CUDA_TRY(cudaMemcpy2DAsync(dst, (size_t)ld * 4, src, (size_t)width * 4, (size_t)width * 4, (size_t)rows,
                                         cudaMemcpyDefault, st));
    }
    return MGGCN_OK;
}

int mggcn_engine_set_inputs(mggcn_engine_t *e, const float *X_host, const int32_t *Y_host, int async) {
    if (!e) return fail(MGGCN_E_INVALID, "null engine");
    TRY(set_dev(e));
    if (async == 2) {
        // prefetch: host -> shadow buffers on the copy stream, concurrent with whatever the compute stream
        // is running; the next epoch / train_forward swaps them in (device-to-device) before it starts
        if (!e->X_shadow) {
            MGGCN_CUDA_TRY(cudaMalloc(&e->X_shadow, (size_t)e->n * (size_t)e->ldX * 4));
            MGGCN_CUDA_TRY(cudaMemset(e->X_shadow, 0, (size_t)e->n * (size_t)e->ldX * 4));
            MGGCN_CUDA_TRY(cudaMalloc(&e->Y_shadow, (size_t)e->n * 4));
            e->X_shadow_alloc = e->X_shadow;
            e->Y_shadow_alloc = e->Y_shadow;
            MGGCN_CUDA_TRY(cudaEventCreateWithFlags(&e->ev_pref, cudaEventDisableTiming));
            MGGCN_CUDA_TRY(cudaEventCreateWithFlags(&e->ev_swap, cudaEventDisableTiming));
        }
        if (e->swap_recorded) MGGCN_CUDA_TRY(cudaStreamWaitEvent(e->scopy, e->ev_swap, 0));  // WAR on the shadows
        if (X_host) TRY(copy_rows_in(e->X_shadow, e->ldX, X_host, e->sizes[0], e->n, e->scopy, e->P));
        if (Y_host) MGGCN_CUDA_TRY(cudaMemcpyAsync(e->Y_shadow, Y_host, (size_t)e->n * 4, cudaMemcpyDefault, e->scopy));
        MGGCN_CUDA_TRY(cudaEventRecord(e->ev_pref, e->scopy));
        e->pref_x = e->pref_x || X_host;
        e->pref_y = e->pref_y || Y_host;
        return MGGCN_OK;
    }
    // ordered after everything already queued on the compute stream (the previous epoch reads X)
    cudaStream_t st = e->sc;
    if (X_host) TRY(copy_rows_in(e->X, e->ldX, X_host, e->sizes[0], e->n, st, e->P));
    if (Y_host) MGGCN_CUDA_TRY(cudaMemcpyAsync(e->Y, Y_host, (size_t)e->n * 4, cudaMemcpyDefault, st));
    if (!async) MGGCN_CUDA_TRY(cudaStreamSynchronize(st));
    return MGGCN_OK;
}

int mggcn_engine_comm_handle(mggcn_engine_t *e, void *out, size_t cap, size_t *len) {
    if (!e || !len) return fail(MGGCN_E_INVALID, "bad argument");
    const size_t need = sizeof(int64_t) * 2 + sizeof(cudaIpcMemHandle_t);
    *len = need;
    if (!out) return MGGCN_OK;
    if (cap < need) return fail(MGGCN_E_INVALID, "handle buffer too small");
    TRY(set_dev(e));
    char *o = static_cast<char *>(out);
    const int64_t pid = (int64_t)getpid();
    const int64_t ptr = (int64_t)reinterpret_cast<uintptr_t>(e->arena);
    memcpy(o, &pid, 8);
    memcpy(o + 8, &ptr, 8);
    cudaIpcMemHandle_t h;
    memset(&h, 0, sizeof h);
    if (e->p2p) MGGCN_CUDA_TRY(cudaIpcGetMemHandle(&h, e->arena));
    memcpy(o + 16, &h, sizeof h);
    return MGGCN_OK;
}

int mggcn_engine_comm_connect(mggcn_engine_t *e, const void *handles, size_t len) {
    if (!e || !handles) return fail(MGGCN_E_INVALID, "bad argument");
    if (!e->p2p) return MGGCN_OK;
    const size_t need = sizeof(int64_t) * 2 + sizeof(cudaIpcMemHandle_t);
    if (len != need) return fail(MGGCN_E_INVALID, "handle length mismatch");
    TRY(set_dev(e));
    const char *hs = static_cast<const char *>(handles);
    for (int r = 0; r < e->P; ++r) {
        if (r == e->rank) continue;
        int64_t pid, ptr;
        memcpy(&pid, hs + (size_t)r * need, 8);
        memcpy(&ptr, hs + (size_t)r * need + 8, 8);
        if (pid == (int64_t)getpid()) {
            // same process (the mg_gcn CLI drives all GPUs from one process): plain peer access
            if (e->fused)
                return fail(MGGCN_E_UNSUPPORTED, "MGGCN_COMM_P2P (one gathered launch per aggregation) needs one process per GPU: "
                                                 "engines that share a process use MGGCN_COMM_P2P_STAGED");
            cudaPointerAttributes attr;
            MGGCN_CUDA_TRY(cudaPointerGetAttributes(&attr, reinterpret_cast<void *>((uintptr_t)ptr)));
            cudaError_t pe = cudaDeviceEnablePeerAccess(attr.device, 0);
            if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled)
                return fail((int)pe, "cudaDeviceEnablePeerAccess(%d): %s", attr.device, cudaGetErrorString(pe));
            cudaGetLastError();
            e->peer_arena[r] = reinterpret_cast<char *>((uintptr_t)ptr);
            e->peer_inproc[r] = true;
        } else {
            cudaIpcMemHandle_t h;
            memcpy(&h, hs + (size_t)r * need + 16, sizeof h);
            void *p = nullptr;
            MGGCN_CUDA_TRY(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
            e->peer_arena[r] = static_cast<char *>(p);
            e->peer_opened[r] = true;
        }
    }
    e->connected = true;
    return MGGCN_OK;
}

int mggcn_engine_train_forward(mggcn_engine_t *e) {
    if (!e) return fail(MGGCN_E_INVALID, "null engine");
    TRY(check_ready(e));
    TRY(set_dev(e));
    TRY(swap_in_inputs(e));
    begin_pass(e);
    TRY(forward(e));
    e->forward_done = true;
    return MGGCN_OK;
}

int mggcn_engine_backward(mggcn_engine_t *e) {
    if (!e) return fail(MGGCN_E_INVALID, "null engine");
    if (!e->forward_done) return fail(MGGCN_E_STATE, "backward before train_forward");
    TRY(set_dev(e));
    TRY(backward(e));
    e->forward_done = false;
    return MGGCN_OK;
}

int mggcn_engine_adam_update(mggcn_engine_t *e, float lr, float beta1, float beta2, float wd, float eps) {
    if (!e) return fail(MGGCN_E_INVALID, "null engine");
    TRY(set_dev(e));
    return adam(e, lr, beta1, beta2, wd, eps);
}

// Captures and instantiates the epoch graph (no launch).  Instantiation and upload allocate device
// memory, which -- with peer access enabled -- waits for the peer GPUs to go idle: a process that
// drives several engines must prepare all of them BEFORE it launches any (a launched engine spins in
// its first barrier until its peers are launched too).
static int ensure_graph(mggcn_engine_t *e, float lr, float beta1, float beta2, float wd, float eps) {
    if ((e->graphs[0] || e->graphs[1]) &&
        (lr != e->g_lr || beta1 != e->g_b1 || beta2 != e->g_b2 || wd != e->g_wd || eps != e->g_eps ||
         e->g_skip != g_skip_exchange || e->g_xs != g_xstreams || e->g_xp != g_xpull))
        drop_graphs(e);
    cudaGraphExec_t &graph_exec = e->graphs[e->flipped ? 1 : 0];  // kernels and tensor maps hold X / Y by address
    if (graph_exec) return MGGCN_OK;
    cudaGraph_t graph = nullptr;
    const uint64_t before = mggcn_launch_count();
    MGGCN_CUDA_TRY(cudaStreamBeginCapture(e->sc, cudaStreamCaptureModeThreadLocal));
    e->capturing = true;
    int rc = epoch_body(e, lr, beta1, beta2, wd, eps);
    e->capturing = false;
    if (rc == MGGCN_OK && e->sm != e->sc) {
        // join the exchange stream back into the origin stream before the capture ends
        cudaError_t ce = cudaEventRecord(e->ev_join, e->sm);
        if (ce == cudaSuccess) ce = cudaStreamWaitEvent(e->sc, e->ev_join, 0);
        if (ce != cudaSuccess) rc = fail((int)ce, "graph join: %s", cudaGetErrorString(ce));
    }
    cudaError_t ce = cudaStreamEndCapture(e->sc, &graph);
    if ((rc || ce != cudaSuccess) && e->fused && g_memop_in_capture && stream_write32()) {
        // this driver does not capture stream memory operations: capture again with memsets as flags
        if (graph) cudaGraphDestroy(graph);
        cudaGetLastError();
        g_memop_in_capture = false;
        g_launches.fetch_sub(mggcn_launch_count() - before, std::memory_order_relaxed);
        return ensure_graph(e, lr, beta1, beta2, wd, eps);
    }
    if (rc) {
        if (graph) cudaGraphDestroy(graph);
        return rc;
    }
    if (ce != cudaSuccess) return fail((int)ce, "cudaStreamEndCapture: %s", cudaGetErrorString(ce));
    e->launches_per_epoch = (int64_t)(mggcn_launch_count() - before);
    g_launches.fetch_sub((uint64_t)e->launches_per_epoch, std::memory_order_relaxed);  // captured, not launched
    ce = cudaGraphInstantiate(&graph_exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ce != cudaSuccess) return fail((int)ce, "cudaGraphInstantiate: %s", cudaGetErrorString(ce));
    MGGCN_CUDA_TRY(cudaGraphUpload(graph_exec, e->sc));
    MGGCN_CUDA_TRY(cudaStreamSynchronize(e->sc));
    e->g_lr = lr; e->g_b1 = beta1; e->g_b2 = beta2; e->g_wd = wd; e->g_eps = eps;
    e->g_skip = g_skip_exchange;
    e->g_xs = g_xstreams;
    e->g_xp = g_xpull;
    return MGGCN_OK;
}

int mggcn_engine_prepare(mggcn_engine_t *e, float lr, float beta1, float beta2, float wd, float eps) {
    if (!e) return fail(MGGCN_E_INVALID, "null engine");
    TRY(check_ready(e));
    TRY(set_dev(e));
    if (!e->use_graph) return MGGCN_OK;
    return ensure_graph(e, lr, beta1, beta2, wd, eps);
}

int mggcn_engine_epoch(mggcn_engine_t *e, float lr, float beta1, float beta2, float wd, float eps) {
    if (!e) return fail(MGGCN_E_INVALID, "null engine");
    TRY(check_ready(e));
    TRY(set_dev(e));
    TRY(swap_in_inputs(e));  // may flip the input buffers: before the graph for the current assignment is looked up
    if (e->use_graph) TRY(ensure_graph(e, lr, beta1, beta2, wd, eps));
    if (!e->use_graph) {
        const uint64_t before = mggcn_launch_count();
        TRY(epoch_body(e, lr, beta1, beta2, wd, eps));
        e->launches_per_epoch = (int64_t)(mggcn_launch_count() - before);
        return MGGCN_OK;
    }
    count_launch((uint64_t)e->launches_per_epoch);
    MGGCN_CUDA_TRY(cudaGraphLaunch(e->graphs[e->flipped ? 1 : 0], e->sc));
    return MGGCN_OK;
}

static int report_dev_err(const mggcn_engine_t *e, const int *err) {
    if (err[0] == 2) return fail(MGGCN_E_STATE, "a feature block of a peer never arrived (stage flag timed out)");
    if (err[0])
        return fail(MGGCN_E_STATE, "peer barrier timed out (a rank is missing or died): rank %d, barrier %d waited for "
                    "rank %d's ticket %d, saw %d", e->rank, err[1] / kMaxRanks, err[1] % kMaxRanks, err[3], err[2]);
    return MGGCN_OK;
}

int mggcn_engine_sync(mggcn_engine_t *e) {
    if (!e) return fail(MGGCN_E_INVALID, "null engine");
    TRY(set_dev(e));
    MGGCN_CUDA_TRY(cudaStreamSynchronize(e->sc));
    if (e->sm != e->sc) MGGCN_CUDA_TRY(cudaStreamSynchronize(e->sm));
    int err[4] = {0, 0, 0, 0};
    MGGCN_CUDA_TRY(cudaMemcpy(err, e->dev_err, sizeof err, cudaMemcpyDefault));
    return report_dev_err(e, err);
}

int mggcn_engine_loss(mggcn_engine_t *e, double *loss_sum, double *correct) {
    if (!e) return fail(MGGCN_E_INVALID, "null engine");
    TRY(set_dev(e));
    double h[2] = {0, 0};
    int err[4] = {0, 0, 0, 0};
    MGGCN_CUDA_TRY(cudaMemcpyAsync(h, e->loss_out, sizeof h, cudaMemcpyDefault, e->sc));
    MGGCN_CUDA_TRY(cudaMemcpyAsync(err, e->dev_err, sizeof err, cudaMemcpyDefault, e->sc));
    MGGCN_CUDA_TRY(cudaStreamSynchronize(e->sc));
    TRY(report_dev_err(e, err));  // a loop that only ever reads the loss must not train on stale peer data silently
    if (loss_sum) *loss_sum = h[0];
    if (correct) *correct = h[1];
    return MGGCN_OK;
}

void *mggcn_engine_stream(mggcn_engine_t *e, int which) {
    if (!e) return nullptr;
    return which == 0 ? e->sc : which == 1 ? e->sm : e->scopy;
}

static int find_tensor(mggcn_engine_t *e, const char *name, float **ptr, int64_t *rows, int64_t *cols, int64_t *ld) {
    if (!name || !name[0]) return fail(MGGCN_E_INVALID, "empty tensor name");
    const bool g = name[0] == 'G';
    const int l = atoi(name + (g ? 2 : 1));
    if (l < 0 || l >= e->L) return fail(MGGCN_E_INVALID, "no layer %d", l);
    layer_t &y = e->layers[l];
    if (name[0] == 'W') { *ptr = y.W; *rows = y.in; *cols = y.out; *ld = y.out; }
    else if (name[0] == 'b') { *ptr = y.b; *rows = 1; *cols = y.out; *ld = y.out; }
    else if (g && name[1] == 'W') { *ptr = y.GW - e->grads + e->gsum; *rows = y.in; *cols = y.out; *ld = y.out; }
    else if (g && name[1] == 'b') { *ptr = y.Gb - e->grads + e->gsum; *rows = 1; *cols = y.out; *ld = y.out; }
    else if (name[0] == 'H') { *ptr = y.Z; *rows = e->n; *cols = y.out; *ld = up4(y.out); }
    else return fail(MGGCN_E_INVALID, "unknown tensor %s", name);
    return MGGCN_OK;
}

int mggcn_engine_get(mggcn_engine_t *e, const char *name, float *out, size_t cap, size_t *count) {
    if (!e || !count) return fail(MGGCN_E_INVALID, "bad argument");
    TRY(set_dev(e));
    float *p; int64_t r, c, ld;
    TRY(find_tensor(e, name, &p, &r, &c, &ld));
    *count = (size_t)(r * c);
    if (!out) return MGGCN_OK;
    if (cap < *count) return fail(MGGCN_E_INVALID, "buffer too small for %s", name);
    TRY(mggcn_engine_sync(e));
    MGGCN_CUDA_TRY(cudaMemcpy2D(out, (size_t)c * 4, p, (size_t)ld * 4, (size_t)c * 4, (size_t)r, cudaMemcpyDefault));
    return MGGCN_OK;
}

int mggcn_engine_set(mggcn_engine_t *e, const char *name, const float *in, size_t count) {
    if (!e || !in) return fail(MGGCN_E_INVALID, "bad argument");
    TRY(set_dev(e));
    float *p; int64_t r, c, ld;
    TRY(find_tensor(e, name, &p, &r, &c, &ld));
    if (count != (size_t)(r * c)) return fail(MGGCN_E_INVALID, "%s has %lld elements", name, (long long)(r * c));
    TRY(mggcn_engine_sync(e));
    MGGCN_CUDA_TRY(cudaMemcpy2D(p, (size_t)ld * 4, in, (size_t)c * 4, (size_t)c * 4, (size_t)r, cudaMemcpyDefault));
    return MGGCN_OK;
}

int mggcn_engine_dump_timers(mggcn_engine_t *e, const char *prefix, char *buf, size_t cap, size_t *written) {
    if (!e || !written) return fail(MGGCN_E_INVALID, "bad argument");
    TRY(mggcn_engine_sync(e));
    std::string s;
    for (const named_timer &t : e->timers) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, t.beg, t.end) != cudaSuccess) { cudaGetLastError(); continue; }
        s += (prefix ? prefix : "");
        s += t.name + ":" + std::to_string(ms) + "\n";  // "<name>:<ms>\n", src/matrix.hpp:149-157
    }
    *written = s.size();
    if (buf && cap) {
        const size_t k = std::min(cap - 1, s.size());
        memcpy(buf, s.data(), k);
        buf[k] = 0;
    }
    return MGGCN_OK;
}

int mggcn_engine_spmm_configs(mggcn_engine_t *e, char *buf, size_t cap, size_t *written) {
    if (!e || !written) return fail(MGGCN_E_INVALID, "bad argument");
    std::string s;
    for (const named_timer &t : e->timers)
        if (!t.cfg.empty()) s += t.name + "\t" + t.cfg + "\n";
    *written = s.size();
    if (buf && cap) {
        const size_t k = std::min(cap - 1, s.size());
        memcpy(buf, s.data(), k);
        buf[k] = 0;
    }
    return MGGCN_OK;
}

int mggcn_engine_get_rows(mggcn_engine_t *e, const char *name, int64_t row_begin, int64_t row_end, float *out, size_t cap,
                          size_t *count) {
    if (!e || !count) return fail(MGGCN_E_INVALID, "bad argument");
    TRY(set_dev(e));
    float *p; int64_t r, c, ld;
    TRY(find_tensor(e, name, &p, &r, &c, &ld));
    if (row_begin < 0 || row_begin > row_end || row_end > r) return fail(MGGCN_E_INVALID, "rows [%lld, %lld) outside %s", (long long)row_begin, (long long)row_end, name);
    *count = (size_t)((row_end - row_begin) * c);
    if (!out) return MGGCN_OK;
    if (cap < *count) return fail(MGGCN_E_INVALID, "buffer too small for %s", name);
    TRY(mggcn_engine_sync(e));
    if (*count)
        MGGCN_CUDA_TRY(cudaMemcpy2D(out, (size_t)c * 4, p + row_begin * ld, (size_t)ld * 4, (size_t)c * 4, (size_t)(row_end - row_begin),
                                    cudaMemcpyDefault));
    return MGGCN_OK;
}

int mggcn_engine_set_mode(mggcn_engine_t *e, int use_graph, int timers) {
    if (!e) return fail(MGGCN_E_INVALID, "null engine");
    TRY(mggcn_engine_sync(e));
    e->use_graph = use_graph != 0;
    e->timers_on = timers != 0 && !e->use_graph;
    return MGGCN_OK;
}

int64_t mggcn_engine_launches_per_epoch(mggcn_engine_t *e) { return e ? e->launches_per_epoch : 0; }

void mggcn_engine_destroy(mggcn_engine_t *e) {
    if (!e) return;
    cudaSetDevice(e->device);
    if (e->sc) cudaStreamSynchronize(e->sc);
    if (e->sm && e->sm != e->sc) cudaStreamSynchronize(e->sm);
    drop_graphs(e);
    for (layer_t &y : e->layers)
        if (y.Wpad) cudaFree(y.Wpad);
    if (e->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(e->comm);
    for (int r = 0; r < e->P; ++r)
        if (e->peer_opened[r]) cudaIpcCloseMemHandle(e->peer_arena[r]);
    for (int w = 0; w < 2; ++w)
        for (int j = 0; j < kMaxRanks; ++j) {
            adj_block &a = e->adj[w][j];
            if (a.plan) mggcn_spmm_plan_destroy(a.plan);
            if (a.rowptr) cudaFree(a.rowptr);
            if (a.colidx) cudaFree(a.colidx);
            if (a.vals) cudaFree(a.vals);
        }
    if (e->ws_spmm) cudaFree(e->ws_spmm);
    if (e->arena) cudaFree(e->arena);
    for (named_timer &t : e->timers) { cudaEventDestroy(t.beg); cudaEventDestroy(t.end); }
    for (cudaEvent_t ev : e->ev_xfer) if (ev) cudaEventDestroy(ev);
    for (cudaEvent_t ev : e->ev_spmm) if (ev) cudaEventDestroy(ev);
    if (e->ev_ready) cudaEventDestroy(e->ev_ready);
    if (e->ev_done) cudaEventDestroy(e->ev_done);
    if (e->ev_join) cudaEventDestroy(e->ev_join);
    if (e->ev_bar) cudaEventDestroy(e->ev_bar);
    if (e->ev_copied) cudaEventDestroy(e->ev_copied);
    if (e->ev_pref) cudaEventDestroy(e->ev_pref);
    if (e->ev_swap) cudaEventDestroy(e->ev_swap);
    if (e->X_shadow_alloc) cudaFree(e->X_shadow_alloc);
    if (e->Y_shadow_alloc) cudaFree(e->Y_shadow_alloc);
    if (e->sm && e->sm != e->sc) cudaStreamDestroy(e->sm);
    if (e->sc) cudaStreamDestroy(e->sc);
    if (e->scopy) cudaStreamDestroy(e->scopy);
    for (int i = 1; i < mggcn_engine::kXStreams; ++i) {
        if (e->sx[i]) cudaStreamDestroy(e->sx[i]);
        if (e->ev_xjoin[i]) cudaEventDestroy(e->ev_xjoin[i]);
    }
    if (e->ev_xfork) cudaEventDestroy(e->ev_xfork);
    cudaGetLastError();
    delete e;
}

int mggcn_bcast_stage_f32(void *nccl_comm, mggcn_stream_t stream, const float *send, float *recv, size_t count,
                          int root) {
    if (!nccl_comm) return fail(MGGCN_E_INVALID, "null communicator");
    TRY(nccl_load());
    NCCL_TRY(g_nccl.Broadcast(send, recv, count, ncclFloat, root, static_cast<ncclComm_t>(nccl_comm),
                              static_cast<cudaStream_t>(stream)));
    return MGGCN_OK;
}

}  // extern "C"
