// CSR x dense SpMM for sm_100a:  C = alpha * A * B + beta * C  (+ fused epilogue).
//
// Replaces matmul(context, csr_matrix, dn_matrix, dn_matrix, ...) = cusparseSpMM(ALG_DEFAULT,
// row-major B/C) of the reference (src/cuda_utils.hpp:27-32) and its workspace query
// (src/cuda_utils.hpp:94-102).
//
// Work decomposition
//   * A work item is (row, [begin, begin+len)) -- a whole row, or one segment of a long row
//     (len <= split_nnz).  The plan (built once per matrix block on the host) lists the segments
//     first and then the whole rows, both longest-first, so the hardware CTA scheduler hands out
//     the heaviest work first and the tail is made of short rows.
//   * One item is processed by a group of LANES threads (LANES = 2..32, chosen from the feature
//     width so that a group covers the row of B with 128-bit loads); a warp holds 32/LANES items.
//     Each lane keeps NCH float4 accumulators in registers.
//   * Column indices and values are read LANES at a time (one coalesced load per lane, the next
//     chunk is prefetched while the current one is consumed) and handed round the group with
//     warp shuffles; the gathers of B rows are issued UNROLL at a time before the FMAs so that
//     UNROLL x 512 B are in flight per warp.
//   * Whole rows are finished in registers: alpha, beta * C, bias, activation / activation
//     gradient / keep-mask, one store.  Segments store their partial sums into the workspace and a
//     second small kernel adds them in segment order (deterministic: no atomics anywhere).
//   * B is the operand with reuse (each row is gathered avg-degree times): it is loaded with an
//     L2 evict_last policy, the CSR stream with evict_first / no L1 allocation.
#include <algorithm>
#include <atomic>
#include <numeric>
#include <vector>

#include "common.hpp"

using namespace mggcn;

namespace mggcn {
extern int g_gemm_tc, g_gemm_kgroup, g_gemm_split, g_gemm_presplit, g_gemm_ts, g_gemm_stages, g_gemm_bn;  // gemm_tc.cu
extern int g_skip_exchange;           // engine.cu
extern int g_xstreams, g_xpull, g_h2d_sm, g_h2d_chunk_kb, g_input_flip, g_xgroups;  // engine.cu
}

namespace {

struct alignas(16) spmm_item {
    uint64_t begin;  // absolute offset into colidx / vals
    uint32_t len;
    uint32_t row;
};
struct split_row {
    uint32_t row, slot0, nseg;
};

}  // namespace

// One launch worth of work: items [first_item, first_item + n_items) of the plan's list; the first n_seg of them
// are segments of split rows (segment k writes partial slot k), rows [first_split, first_split + n_split) of the
// split list are finished by the fix-up kernel.
struct plan_block {
    int64_t first_item = 0, n_items = 0, n_seg = 0, first_split = 0, n_split = 0;
    int64_t nnz = 0;  // entries covered by the items
};

struct mggcn_spmm_plan {
    int64_t rows = 0;
    int split_nnz = 0;
    int device = 0;
    // blk.size() == 1: the whole matrix in one launch.  > 1: column blocks (2-D schedule, see plan_create_blocked):
    // block j holds, per row, the entries whose column falls into the j-th range of `blk_cols` columns; the blocks
    // are multiplied one after the other (and per feature window) so that the rows of B a launch gathers stay in L2.
    std::vector<plan_block> blk;
    int64_t blk_cols = 0;   // columns per block (the last one may be smaller)
    int64_t max_seg = 0;    // max over blocks of n_seg (workspace slots)
    spmm_item *d_items = nullptr;
    split_row *d_split = nullptr;
};

namespace {

int g_unroll = 4;     // gathers in flight per lane (4 or 8)
int g_threads = 64;   // CTA size
int g_policy = 1;     // 1: L2 evict_last for B, evict_first for the CSR stream; 0: plain loads
int g_split = 0;      // row-split threshold of new plans; 0 = chosen from the matrix (see plan_create)
int g_slices = 0;     // k > 1: run as k launches over F/k-wide column windows of B and C; 0: k = 2 when B (cols x F floats)
                      // is larger than half of the L2 but its half is not (see spmm_scheduled); 1: never
int g_narrow = 0;     // lane mapping for 9..16 float4 per row, see spmm_dispatch
int g_tma = 0;        // 1: F = 64 / 128 products gather B through the TMA unit (spmm_tma_kernel, experiment)
int g_blocks = 0;     // column blocks of new blocked plans: 0 / 1 = none (measured best), k > 1 = exactly k, < 0 = traffic model
int g_window = 0;     // feature-window width of blocked plans at launch: 0 = widest that keeps a block's B rows resident
int g_tile_mb = 0;    // bytes of B a launch may count on staying L2-resident, in MiB; 0 = 45 % of the L2

struct spmm_args {
    const spmm_item *items;  // null: natural order, one item per row from rowptr
    int64_t n_items;
    int64_t n_seg;
    const void *rowptr;
    int rowptr64;
    const uint32_t *colidx;
    const float *vals;
    const float *B;
    int64_t ldb;
    float *C;
    int64_t ldc;
    float *partial;
    int64_t ldp;
    int F;
    float alpha, beta;
    mggcn_epilogue epi;
    int policy;
    // "gathered" multi-GPU form (GATHER kernels only): B is the all-gather buffer in global row order that
    // the copy engines are still filling while the kernel runs; this rank's own rows (columns
    // [own_lo, own_lo + own_n)) are copied into it on the launching stream right before the kernel, so every
    // gather is one address computation.  flags[s] != 0 once the block of stage s has landed.
    mggcn::spmm_gather g;
};

__device__ __forceinline__ uint64_t make_policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t make_policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t make_policy_normal() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ float4 ldg_f4(const float *p, uint64_t pol) {
    float4 r;
    asm volatile("ld.global.nc.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p), "l"(pol));
    return r;
}
__device__ __forceinline__ float ldg_f1(const float *p, uint64_t pol) {
    float r;
    asm volatile("ld.global.nc.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(r) : "l"(p), "l"(pol));
    return r;
}
__device__ __forceinline__ uint32_t ldg_stream_u32(const uint32_t *p, uint64_t pol) {
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.u32 %0, [%1], %2;"
                 : "=r"(r)
                 : "l"(p), "l"(pol));
    return r;
}
__device__ __forceinline__ float ldg_stream_f32(const float *p, uint64_t pol) {
    float r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.f32 %0, [%1], %2;"
                 : "=f"(r)
                 : "l"(p), "l"(pol));
    return r;
}

__device__ __forceinline__ float4 ldg_coherent_f4(const float *p, uint64_t pol) {
    float4 r;
    asm volatile("ld.global.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p), "l"(pol)
                 : "memory");
    return r;
}
__device__ __forceinline__ float ldg_coherent_f1(const float *p, uint64_t pol) {
    float r;
    asm volatile("ld.global.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(r) : "l"(p), "l"(pol) : "memory");
    return r;
}
// Spins until the copy engine has published stage `s` (flags are written after the block's data on
// the same stream).  Bounded: about 4 s of SM clock, then the error word is raised and the kernel
// carries on with whatever is there, so a wiring mistake cannot hang the GPU.
__device__ __forceinline__ void wait_stage(const uint32_t *flag, int *err) {
    uint32_t v;
    const long long t0 = clock64();
    for (;;) {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
        if (v) break;
        if (clock64() - t0 > 8000000000ll) {
            if (err) *err = 2;
            break;
        }
        __nanosleep(64);
    }
}

template <int VEC>
struct vec_t;
template <>
struct vec_t<4> {
    float v[4];
    __device__ __forceinline__ void load(const float *p, uint64_t pol) {
        float4 t = ldg_f4(p, pol);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
    __device__ __forceinline__ void load_coherent(const float *p, uint64_t pol) {
        float4 t = ldg_coherent_f4(p, pol);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
};
template <>
struct vec_t<1> {
    float v[1];
    __device__ __forceinline__ void load(const float *p, uint64_t pol) { v[0] = ldg_f1(p, pol); }
    __device__ __forceinline__ void load_coherent(const float *p, uint64_t pol) { v[0] = ldg_coherent_f1(p, pol); }
};

__device__ __forceinline__ float apply_epilogue(float x, const mggcn_epilogue &e, int64_t row, int col) {
    if (e.bias) x += __ldg(e.bias + col);
    if (e.kind == MGGCN_EPI_LEAKY_RELU) {
        const float ax = e.slope * x;  // max(x, slope*x), src/cuda_utils.cu:26-31
        x = x > ax ? x : ax;
    } else if (e.kind == MGGCN_EPI_LEAKY_RELU_BWD) {
        const float z = __ldg(e.aux + row * e.ld_aux + col);  // src/cuda_utils.cu:33-38
        x = z > 0.f ? x : e.slope * x;
    }
    if (e.keep_mask) x = e.keep_mask[row * e.ld_mask + col] ? x * e.keep_scale : 0.f;
    return x;
}

// One group of LANES threads per item; NCH vectors of VEC floats per lane; the group covers
// LANES*NCH*VEC features per pass and loops over passes when F is wider.
// EXACT: F is a multiple of the pass width, so no per-lane column predicate is needed.
template <int LANES, int NCH, int VEC, int UNROLL, bool EXACT, bool GATHER>
__device__ __forceinline__ void spmm_body(const spmm_args &a) {
    constexpr int GROUPS = 32 / LANES;
    constexpr int TILE_F = LANES * NCH * VEC;
    const int lane = threadIdx.x & 31;
    const int sl = lane & (LANES - 1);
    const int sub = lane / LANES;
    const int warps_per_cta = blockDim.x >> 5;
    const int64_t warp_id = (int64_t)blockIdx.x * warps_per_cta + (threadIdx.x >> 5);
    const int64_t item_idx = warp_id * GROUPS + sub;
    const bool valid = item_idx < a.n_items;

    uint64_t begin = 0;
    uint32_t len = 0, row = 0;
    if (valid) {
        if (a.items) {
            const spmm_item it = a.items[item_idx];
            begin = it.begin; len = it.len; row = it.row;
        } else {
            row = (uint32_t)item_idx;
            if (a.rowptr64) {
                const uint64_t *rp = static_cast<const uint64_t *>(a.rowptr);
                begin = rp[row];
                len = (uint32_t)(rp[row + 1] - begin);
            } else {
                const uint32_t *rp = static_cast<const uint32_t *>(a.rowptr);
                begin = rp[row];
                len = rp[row + 1] - (uint32_t)begin;
            }
        }
    }
    // warp-uniform trip count (groups in one warp hold items of similar length: the plan sorts)
    uint32_t maxlen = len;
    if (GROUPS > 1) {
#pragma unroll
        for (int o = 16; o >= LANES; o >>= 1) maxlen = max(maxlen, __shfl_xor_sync(0xffffffffu, maxlen, o));
    }
    const uint64_t pol_b = a.policy ? make_policy_evict_last() : make_policy_normal();
    const uint64_t pol_s = a.policy ? make_policy_evict_first() : make_policy_normal();
    const uint32_t *ci = a.colidx + begin;
    const float *va = a.vals + begin;
    const bool is_segment = valid && item_idx < a.n_seg;

    for (int f0 = 0; f0 < a.F; f0 += TILE_F) {
        float acc[NCH][VEC];
        bool colok[NCH];
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
            colok[k] = EXACT || f0 + VEC * (sl + LANES * k) < a.F;
#pragma unroll
            for (int e = 0; e < VEC; ++e) acc[k][e] = 0.f;
        }
        const float *Bl = a.B + f0 + VEC * sl;

        uint32_t c_cur = GATHER ? a.g.own_lo : 0, c_nxt = 0;
        float v_cur = 0.f, v_nxt = 0.f;
        // GATHER: the blocks of stages [0, ready] are in place; they cover the cyclic column interval
        // [own_lo, own_lo + g_len) of the gather buffer (own block first, then blocks rank+1, ... cyclically)
        int ready = 0;
        uint32_t g_len = GATHER ? a.g.own_n : 0;
        if ((uint32_t)sl < len) {
            c_cur = ldg_stream_u32(ci + sl, pol_s);
            v_cur = ldg_stream_f32(va + sl, pol_s);
        }
        for (uint32_t pos = 0; pos < maxlen; pos += LANES) {
            const uint32_t npos = pos + LANES;
            if (npos + sl < len) {  // prefetch the next chunk of (column, value) pairs
                c_nxt = ldg_stream_u32(ci + npos + sl, pol_s);
                v_nxt = ldg_stream_f32(va + npos + sl, pol_s);
            } else {
                c_nxt = GATHER ? a.g.own_lo : 0; v_nxt = 0.f;
            }
            if (GATHER) {
                // rows are stored in stage order, so the stage of a chunk only grows: wait (once per stage and
                // warp) until every column this chunk gathers from lies inside the interval that has landed
                uint32_t cyc = c_cur - a.g.own_lo;
                if (c_cur < a.g.own_lo) cyc += a.g.ncols;
                while (__any_sync(0xffffffffu, cyc >= g_len)) {
                    if (++ready >= a.g.P) break;  // malformed input: columns outside every block
                    wait_stage(a.g.flags + ready, a.g.err);
                    int b = a.g.rank + ready;
                    b -= b >= a.g.P ? a.g.P : 0;
                    uint32_t lo = 0, hi = a.g.ncols;
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        if (i == b) lo = a.g.cuts[i];
                        if (i == b + 1 && i < a.g.P) hi = a.g.cuts[i];
                    }
                    g_len += hi - lo;
                }
            }
            const int rem = (int)len - (int)pos;  // entries of this group's item left (may be <= 0)
            const bool full = __all_sync(0xffffffffu, rem >= LANES);
            if (full) {
#pragma unroll
                for (int j = 0; j < LANES; j += UNROLL) {
                    vec_t<VEC> b[UNROLL < LANES ? UNROLL : LANES][NCH];
#pragma unroll
                    for (int u = 0; u < UNROLL && u < LANES; ++u) {
                        const uint32_t c = __shfl_sync(0xffffffffu, c_cur, j + u, LANES);
                        const float *bp = Bl + (int64_t)c * a.ldb;
#pragma unroll
                        for (int k = 0; k < NCH; ++k) {
                            if (EXACT || colok[k]) {
                                if (GATHER) b[u][k].load_coherent(bp + VEC * LANES * k, pol_b);
                                else b[u][k].load(bp + VEC * LANES * k, pol_b);
                            } else {
#pragma unroll
                                for (int e = 0; e < VEC; ++e) b[u][k].v[e] = 0.f;
                            }
                        }
                    }
#pragma unroll
                    for (int u = 0; u < UNROLL && u < LANES; ++u) {
                        const float v = __shfl_sync(0xffffffffu, v_cur, j + u, LANES);
#pragma unroll
                        for (int k = 0; k < NCH; ++k)
#pragma unroll
                            for (int e = 0; e < VEC; ++e) acc[k][e] = fmaf(v, b[u][k].v[e], acc[k][e]);
                    }
                }
            } else {
                int maxrem = rem;
                if (GROUPS > 1) {
#pragma unroll
                    for (int o = 16; o >= LANES; o >>= 1) maxrem = max(maxrem, __shfl_xor_sync(0xffffffffu, maxrem, o));
                }
                for (int j = 0; j < LANES && j < maxrem; j += UNROLL) {
                    vec_t<VEC> b[UNROLL < LANES ? UNROLL : LANES][NCH];
                    float v[UNROLL < LANES ? UNROLL : LANES];
#pragma unroll
                    for (int u = 0; u < UNROLL && u < LANES; ++u) {
                        const uint32_t c = __shfl_sync(0xffffffffu, c_cur, j + u, LANES);
                        v[u] = __shfl_sync(0xffffffffu, v_cur, j + u, LANES);
                        const bool ok = j + u < rem;
                        const float *bp = Bl + (int64_t)c * a.ldb;
#pragma unroll
                        for (int k = 0; k < NCH; ++k) {
                            if (ok && colok[k]) {
                                if (GATHER) b[u][k].load_coherent(bp + VEC * LANES * k, pol_b);
                                else b[u][k].load(bp + VEC * LANES * k, pol_b);
                            } else {
#pragma unroll
                                for (int e = 0; e < VEC; ++e) b[u][k].v[e] = 0.f;
                            }
                        }
                        if (!ok) v[u] = 0.f;
                    }
#pragma unroll
                    for (int u = 0; u < UNROLL && u < LANES; ++u)
#pragma unroll
                        for (int k = 0; k < NCH; ++k)
#pragma unroll
                            for (int e = 0; e < VEC; ++e) acc[k][e] = fmaf(v[u], b[u][k].v[e], acc[k][e]);
                }
            }
            c_cur = c_nxt; v_cur = v_nxt;
        }

        if (!valid) continue;
        if (is_segment) {  // partial sums of a long row's segment; combined by spmm_fixup_kernel
            float *p = a.partial + item_idx * a.ldp + f0 + VEC * sl;
#pragma unroll
            for (int k = 0; k < NCH; ++k) {
                if (!colok[k]) continue;
                if constexpr (VEC == 4) {
                    *reinterpret_cast<float4 *>(p + 4 * LANES * k) =
                        make_float4(acc[k][0], acc[k][1], acc[k][2], acc[k][3]);
                } else {
                    p[LANES * k] = acc[k][0];
                }
            }
            continue;
        }
        float *crow = a.C + (int64_t)row * a.ldc;
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
            if (!colok[k]) continue;
            const int col0 = f0 + VEC * (sl + LANES * k);
            if constexpr (VEC == 4) {
                if (EXACT || col0 + 3 < a.F) {
                    float out[4];
                    float4 old = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (a.beta != 0.f) old = *reinterpret_cast<const float4 *>(crow + col0);
                    out[0] = a.alpha * acc[k][0] + (a.beta != 0.f ? a.beta * old.x : 0.f);
                    out[1] = a.alpha * acc[k][1] + (a.beta != 0.f ? a.beta * old.y : 0.f);
                    out[2] = a.alpha * acc[k][2] + (a.beta != 0.f ? a.beta * old.z : 0.f);
                    out[3] = a.alpha * acc[k][3] + (a.beta != 0.f ? a.beta * old.w : 0.f);
#pragma unroll
                    for (int e = 0; e < 4; ++e) out[e] = apply_epilogue(out[e], a.epi, row, col0 + e);
                    *reinterpret_cast<float4 *>(crow + col0) = make_float4(out[0], out[1], out[2], out[3]);
                    continue;
                }
            }
            {
#pragma unroll
                for (int e = 0; e < VEC; ++e) {
                    if (col0 + e >= a.F) break;
                    float x = a.alpha * acc[k][e];
                    if (a.beta != 0.f) x += a.beta * crow[col0 + e];
                    crow[col0 + e] = apply_epilogue(x, a.epi, row, col0 + e);
                }
            }
        }
    }
}

template <int LANES, int NCH, int VEC, int UNROLL, bool EXACT, bool GATHER = false>
__global__ void __launch_bounds__(256) spmm_kernel(const spmm_args a) {
    spmm_body<LANES, NCH, VEC, UNROLL, EXACT, GATHER>(a);
}
// Adds the partial sums of every split row in segment order and finishes the row.
__global__ void spmm_fixup_kernel(const split_row *rows, int64_t n_split, const float *partial,
                                  int64_t ldp, float *C, int64_t ldc, int F, float alpha, float beta,
                                  const mggcn_epilogue epi) {
    const int64_t total = n_split * (int64_t)F;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t s = i / F;
        const int col = (int)(i - s * F);
        const split_row r = rows[s];
        float acc = 0.f;
        for (uint32_t k = 0; k < r.nseg; ++k) acc += partial[(int64_t)(r.slot0 + k) * ldp + col];
        float *c = C + (int64_t)r.row * ldc + col;
        float x = alpha * acc;
        if (beta != 0.f) x += beta * *c;
        *c = apply_epilogue(x, epi, r.row, col);
    }
}

// ---------------------------------------------------------------------------------------------
// Experiment (mggcn_tune("tma", 1)): the rows of B travel L2 -> shared memory through the
// TMA unit (cp.async.bulk, one 1-D bulk copy per gathered row, mbarrier-paced) instead of LSU loads into
// registers, so the L1TEX pipe that ncu shows at 88 % carries none of the gather stream.  One warp per item; a ring
// of TS stages x TR rows per warp; lanes 0..TR-1 issue the copies of a stage (lane j the row of entry j), every lane
// then reads its float4 of each landed row from shared memory (conflict-free: consecutive lanes, consecutive 16 B).
// Exact widths only (F = 4 * LPR floats, LPR = 16 or 32 lanes per row; with LPR = 16 the two half-warps take
// alternate rows and are added at the end).  Result: see DESIGN.md section 7 -- kept or dropped on the measurement.
// ---------------------------------------------------------------------------------------------
constexpr int TR = 8, TS = 4;

__device__ __forceinline__ uint32_t tma_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int LPR>
__global__ void __launch_bounds__(128) spmm_tma_kernel(const spmm_args a) {
    constexpr int RB = LPR * 16;      // bytes per row of B
    constexpr int RP = 32 / LPR;      // rows a warp reads per shared-memory instruction
    extern __shared__ __align__(128) uint8_t tma_smem[];
    const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5, nw = blockDim.x >> 5;
    uint8_t *ring = tma_smem + (size_t)wic * (TS * TR * RB);
    const uint32_t ring_u32 = tma_smem_u32(ring);
    const uint32_t bar0 = tma_smem_u32(tma_smem + (size_t)nw * (TS * TR * RB)) + (uint32_t)wic * TS * 8;
    if (lane < TS) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8u * lane));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();

    const int64_t item_idx = (int64_t)blockIdx.x * nw + wic;
    if (item_idx >= a.n_items) return;
    uint64_t begin;
    uint32_t len, row;
    if (a.items) {
        const spmm_item it = a.items[item_idx];
        begin = it.begin; len = it.len; row = it.row;
    } else {
        row = (uint32_t)item_idx;
        const uint32_t *rp = static_cast<const uint32_t *>(a.rowptr);
        begin = rp[row];
        len = rp[row + 1] - (uint32_t)begin;
    }
    const uint64_t pol_s = make_policy_evict_first();
    const uint32_t *ci = a.colidx + begin;
    const float *va = a.vals + begin;
    // three register chunks of 32 (column, value) pairs: k (being consumed), k+1 (being issued), k+2 (in flight)
    uint32_t c0 = 0, c1 = 0, c2 = 0;
    float v0 = 0.f, v1 = 0.f, v2 = 0.f;
    auto load_chunk = [&](uint32_t k, uint32_t &c, float &v) {
        const uint32_t e = 32u * k + lane;
        c = 0; v = 0.f;
        if (e < len) {
            c = ldg_stream_u32(ci + e, pol_s);
            v = ldg_stream_f32(va + e, pol_s);
        }
    };
    load_chunk(0, c0, v0);
    load_chunk(1, c1, v1);
    load_chunk(2, c2, v2);
    uint32_t kbase = 0;  // chunk held in c0 / v0
    const uint32_t n_sub = (len + TR - 1) / TR;

    auto issue = [&](uint32_t q) {
        const uint32_t e0 = q * TR, stage = q % TS;
        const uint32_t nvalid = min((uint32_t)TR, len - e0);
        const uint32_t which = e0 / 32 - kbase;  // 0 or 1 (the issue pointer runs (TS-1)*TR = 24 entries ahead)
        const uint32_t c = __shfl_sync(0xffffffffu, which == 0 ? c0 : (which == 1 ? c1 : c2), (e0 + lane) & 31);
        const uint32_t bar = bar0 + 8u * stage;
        if (lane == 0)
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(nvalid * (uint32_t)RB) : "memory");
        __syncwarp();
        if ((uint32_t)lane < nvalid) {
            const float *src = a.B + (int64_t)c * a.ldb;
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(ring_u32 + (stage * TR + lane) * RB), "l"(src), "r"((uint32_t)RB), "r"(bar)
                         : "memory");
        }
    };

    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (uint32_t q = 0; q < (uint32_t)(TS - 1) && q < n_sub; ++q) issue(q);
    for (uint32_t q = 0; q < n_sub; ++q) {
        if (q + TS - 1 < n_sub) issue(q + TS - 1);
        const uint32_t stage = q % TS, parity = (q / TS) & 1u, bar = bar0 + 8u * stage;
        uint32_t done = 0;
        while (!done) {
            asm volatile(
                "{\n\t.reg .pred p;\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                "selp.u32 %0, 1, 0, p;\n\t}"
                : "=r"(done)
                : "r"(bar), "r"(parity)
                : "memory");
        }
        const uint32_t e0 = q * TR;
        const uint32_t which = e0 / 32 - kbase;
        const float4 *rows = reinterpret_cast<const float4 *>(ring + (size_t)stage * TR * RB);
        float4 x[TR / RP];
#pragma unroll
        for (int j = 0; j < TR / RP; ++j) x[j] = rows[(j * RP + lane / LPR) * LPR + (lane % LPR)];
        __syncwarp();  // every lane has read the stage: the next issue may overwrite it
#pragma unroll
        for (int j = 0; j < TR / RP; ++j) {
            const uint32_t e = e0 + j * RP + lane / LPR;
            float v = __shfl_sync(0xffffffffu, which == 0 ? v0 : v1, e & 31);
            if (e < len) {
                acc[0] = fmaf(v, x[j].x, acc[0]); acc[1] = fmaf(v, x[j].y, acc[1]);
                acc[2] = fmaf(v, x[j].z, acc[2]); acc[3] = fmaf(v, x[j].w, acc[3]);
            }
        }
        if (((e0 + TR) & 31u) == 0) {  // the consumer leaves chunk kbase: rotate and prefetch chunk kbase + 3
            c0 = c1; v0 = v1; c1 = c2; v1 = v2;
            ++kbase;
            load_chunk(kbase + 2, c2, v2);
        }
    }
    if (RP == 2) {
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], 16);
        if (lane >= 16) return;
    }
    const int col0 = 4 * (lane % LPR);
    if (item_idx < a.n_seg) {
        *reinterpret_cast<float4 *>(a.partial + item_idx * a.ldp + col0) = make_float4(acc[0], acc[1], acc[2], acc[3]);
        return;
    }
    float *crow = a.C + (int64_t)row * a.ldc;
    float4 old = make_float4(0.f, 0.f, 0.f, 0.f);
    if (a.beta != 0.f) old = *reinterpret_cast<const float4 *>(crow + col0);
    float out[4] = {a.alpha * acc[0] + (a.beta != 0.f ? a.beta * old.x : 0.f), a.alpha * acc[1] + (a.beta != 0.f ? a.beta * old.y : 0.f),
                    a.alpha * acc[2] + (a.beta != 0.f ? a.beta * old.z : 0.f), a.alpha * acc[3] + (a.beta != 0.f ? a.beta * old.w : 0.f)};
#pragma unroll
    for (int e = 0; e < 4; ++e) out[e] = apply_epilogue(out[e], a.epi, row, col0 + e);
    *reinterpret_cast<float4 *>(crow + col0) = make_float4(out[0], out[1], out[2], out[3]);
}

thread_local char t_last_config[256] = "";

template <int LPR>
int launch_tma(cudaStream_t st, const spmm_args &a) {
    constexpr int threads = 128, nw = threads / 32;
    constexpr size_t smem = (size_t)nw * TS * TR * LPR * 16 + (size_t)nw * TS * 8;
    static std::atomic<bool> attr_set[64];
    int dev = 0;
    MGGCN_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev].load(std::memory_order_acquire)) {
        MGGCN_CUDA_TRY(cudaFuncSetAttribute(spmm_tma_kernel<LPR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (dev >= 0 && dev < 64) attr_set[dev].store(true, std::memory_order_release);
    }
    const int64_t grid = (a.n_items + nw - 1) / nw;
    if (grid <= 0) return MGGCN_OK;
    if (grid > 0x7fffffffll) return fail(MGGCN_E_UNSUPPORTED, "SpMM grid too large");
    spmm_tma_kernel<LPR><<<(unsigned)grid, threads, smem, st>>>(a);
    MGGCN_LAUNCH_CHECK();
    snprintf(t_last_config, sizeof t_last_config, "spmm_tma_kernel<%d> (rows of B through cp.async.bulk, %d stages x %d rows per warp) grid=%u cta=%d F=%d",
             LPR, TS, TR, (unsigned)grid, threads, a.F);
    return MGGCN_OK;
}

template <int LANES, int NCH, int VEC>
int launch_cfg(cudaStream_t st, const spmm_args &a) {
    constexpr int GROUPS = 32 / LANES;
    const int threads = g_threads;
    const int64_t items_per_cta = (int64_t)(threads / 32) * GROUPS;
    const int64_t grid = (a.n_items + items_per_cta - 1) / items_per_cta;
    if (grid <= 0) return MGGCN_OK;
    if (grid > 0x7fffffffll) return fail(MGGCN_E_UNSUPPORTED, "SpMM grid too large");
    constexpr int U4 = LANES < 4 ? LANES : 4;
    const bool exact = a.F % (LANES * NCH * VEC) == 0;
    const unsigned gr = (unsigned)grid;
    int unroll = U4;
    if (a.g.flags) {  // gathered multi-GPU operand
        if (exact) spmm_kernel<LANES, NCH, VEC, U4, true, true><<<gr, threads, 0, st>>>(a);
        else spmm_kernel<LANES, NCH, VEC, U4, false, true><<<gr, threads, 0, st>>>(a);
    } else if (g_unroll >= 8 && LANES >= 8) {
        unroll = 8;
        if (exact) spmm_kernel<LANES, NCH, VEC, 8, true><<<gr, threads, 0, st>>>(a);
        else spmm_kernel<LANES, NCH, VEC, 8, false><<<gr, threads, 0, st>>>(a);
    } else {
        if (exact) spmm_kernel<LANES, NCH, VEC, U4, true><<<gr, threads, 0, st>>>(a);
        else spmm_kernel<LANES, NCH, VEC, U4, false><<<gr, threads, 0, st>>>(a);
    }
    MGGCN_LAUNCH_CHECK();
    snprintf(t_last_config, sizeof t_last_config, "spmm_kernel<%d,%d,%d,%d,%s%s> grid=%u cta=%d F=%d", LANES, NCH, VEC, unroll,
             exact ? "exact" : "ragged", a.g.flags ? ",gathered" : "", gr, threads, a.F);
    return MGGCN_OK;
}

// One launch (+ fix-up of its split rows): the whole matrix (plan == null: natural row order), or block `pb` of a plan.
int spmm_dispatch(cudaStream_t st, int64_t rows, int64_t cols, const void *rowptr, int rowptr64,
                  const uint32_t *colidx, const float *vals, const float *B, int64_t ldb, float *C,
                  int64_t ldc, int F, float alpha, float beta, const mggcn_epilogue *epi,
                  const mggcn_spmm_plan_t *plan, const plan_block *pb, void *ws,
                  const mggcn::spmm_gather *gather = nullptr) {
    if (rows < 0 || cols < 0 || F < 0) return fail(MGGCN_E_INVALID, "negative dimension");
    if (rows == 0 || F == 0) return MGGCN_OK;
    if (!rowptr && !plan) return fail(MGGCN_E_INVALID, "rowptr is null");
    if (!C || (!B && cols > 0)) return fail(MGGCN_E_INVALID, "null operand");
    if (ldb < F || ldc < F) return fail(MGGCN_E_INVALID, "leading dimension smaller than F");
    if (rows > 0xFFFFFFFFll) return fail(MGGCN_E_UNSUPPORTED, "more than 2^32 rows in one block");
    if (plan && plan->rows != rows) return fail(MGGCN_E_INVALID, "plan was built for %lld rows, got %lld",
                                               (long long)plan->rows, (long long)rows);
    if (plan && !pb) pb = &plan->blk[0];
    spmm_args a{};
    a.items = plan ? plan->d_items + pb->first_item : nullptr;
    a.n_items = plan ? pb->n_items : rows;
    a.n_seg = plan ? pb->n_seg : 0;
    a.rowptr = rowptr;
    a.rowptr64 = rowptr64;
    a.colidx = colidx;
    a.vals = vals;
    a.B = B;
    a.ldb = ldb;
    a.C = C;
    a.ldc = ldc;
    a.F = F;
    a.alpha = alpha;
    a.beta = beta;
    a.policy = g_policy;
    if (gather) {
        if (!gather->flags || gather->P < 1 || gather->P > 16 || gather->rank < 0 || gather->rank >= gather->P)
            return fail(MGGCN_E_INVALID, "bad gathered-operand description");
        if ((uint64_t)gather->own_lo + gather->own_n > (uint64_t)cols || cols > 0xFFFFFFFFll)
            return fail(MGGCN_E_INVALID, "own column range outside the gathered operand");
        a.g = *gather;
        a.g.ncols = (uint32_t)cols;
        // the own block (stage 0) goes into its place of the gather buffer on this stream, ahead of the kernel
        float *slot = const_cast<float *>(B) + (size_t)gather->own_lo * (size_t)ldb;
        if (gather->B_own && gather->B_own != slot && gather->own_n > 0)
            MGGCN_CUDA_TRY(cudaMemcpyAsync(slot, gather->B_own, (size_t)gather->own_n * (size_t)ldb * sizeof(float),
                                           cudaMemcpyDeviceToDevice, st));
    }
    if (epi) a.epi = *epi;
    else memset(&a.epi, 0, sizeof a.epi);
    if (a.epi.kind == MGGCN_EPI_LEAKY_RELU_BWD && !a.epi.aux)
        return fail(MGGCN_E_INVALID, "LEAKY_RELU_BWD epilogue needs aux");
    const int F4 = (F + 3) / 4;
    a.ldp = (int64_t)F4 * 4;
    a.partial = static_cast<float *>(ws);
    if (a.n_seg > 0 && !ws) return fail(MGGCN_E_INVALID, "plan has split rows but workspace is null");

    const bool aligned = ldb % 4 == 0 && ldc % 4 == 0 && ldb >= 4 * (int64_t)F4 &&
                         (reinterpret_cast<uintptr_t>(B) & 15) == 0 &&
                         (!gather || (reinterpret_cast<uintptr_t>(gather->B_own) & 15) == 0) &&
                         (reinterpret_cast<uintptr_t>(C) & 15) == 0 &&
                         (reinterpret_cast<uintptr_t>(ws) & 15) == 0;
    int rc;
    if (g_tma && aligned && !gather && !rowptr64 && (F == 128 || F == 64) && (ldb * 4) % 16 == 0) {
        rc = F == 128 ? launch_tma<32>(st, a) : launch_tma<16>(st, a);
    } else if (aligned) {
        if (F4 <= 2) rc = launch_cfg<2, 1, 4>(st, a);
        else if (F4 <= 4) rc = launch_cfg<4, 1, 4>(st, a);
        else if (F4 <= 8) rc = launch_cfg<8, 1, 4>(st, a);
        // class widths (41 -> 44, 47 -> 48 floats = 11 / 12 float4): 16 lanes leave 5 / 4 of them idle on every
        // gather; "narrow" = 1: 4 lanes x 3 float4 (8 rows per warp, no idle lane), 2: 8 lanes x 2 float4
        else if (F4 > 8 && F4 <= 12 && g_narrow == 1) rc = launch_cfg<4, 3, 4>(st, a);
        else if (F4 > 8 && F4 <= 16 && g_narrow == 2) rc = launch_cfg<8, 2, 4>(st, a);
        else if (F4 <= 16) rc = launch_cfg<16, 1, 4>(st, a);
        else if (F4 <= 32) rc = launch_cfg<32, 1, 4>(st, a);
        else rc = launch_cfg<32, 2, 4>(st, a);
    } else {
        if (F <= 8) rc = launch_cfg<8, 1, 1>(st, a);
        else if (F <= 16) rc = launch_cfg<16, 1, 1>(st, a);
        else if (F <= 32) rc = launch_cfg<32, 1, 1>(st, a);
        else if (F <= 64) rc = launch_cfg<32, 2, 1>(st, a);
        else rc = launch_cfg<32, 4, 1>(st, a);
    }
    if (rc) return rc;
    if (plan && pb->n_split > 0) {
        const int64_t total = pb->n_split * (int64_t)F;
        const int grid = (int)std::min<int64_t>((total + 255) / 256, 148 * 8);
        spmm_fixup_kernel<<<grid, 256, 0, st>>>(plan->d_split + pb->first_split, pb->n_split, a.partial, a.ldp, C,
                                                ldc, F, alpha, beta, a.epi);
        MGGCN_LAUNCH_CHECK();
    }
    return MGGCN_OK;
}

thread_local int64_t t_touched_cols = -1;

// What a launch can keep resident of the operand with reuse (B): the two L2 halves of the chip each cache the
// lines the SMs of their die ask for, so a B that every SM gathers from stays resident only up to about half of
// the L2 (measured on B200: 60 MB windows hit 88 %, a 119 MB B is re-read from DRAM 9 times over).
int64_t l2_bytes() {
    static int64_t v = [] {
        int dev = 0, l2 = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, dev) != cudaSuccess || l2 <= 0) {
            cudaGetLastError();
            l2 = 126 << 20;
        }
        return (int64_t)l2;
    }();
    return v;
}
int64_t resident_budget() { return g_tile_mb > 0 ? (int64_t)g_tile_mb << 20 : l2_bytes() * 45 / 100; }

// The schedule of one product: feature windows x column blocks.
//   * Column blocks come from the plan (plan_create_blocked): launches walk block after block, C accumulates with
//     beta = 1 between them, the caller's beta belongs to the first block and the epilogue to the last (their item
//     lists cover every row; the blocks in between list only rows that have entries there).
//   * Feature windows: w floats of every row of B and C per pass, chosen so that (columns of a block) x w x 4 bytes
//     stays L2-resident.  Without column blocks this is the feature slicing of round 1: two windows when B is between
//     ~0.7 and 1 L2 (reddit shape, F = 128: 4.4 GB of DRAM traffic per aggregation instead of 10.7).
// The sums per output element are taken block by block in column order: deterministic, and equal to the unblocked
// order up to fp32 rounding of the partial sums (inside the 1e-5 bar, tests/test_gpu_spmm.py).
int spmm_scheduled(cudaStream_t st, int64_t rows, int64_t cols, const void *rowptr, int rowptr64, const uint32_t *colidx,
                   const float *vals, const float *B, int64_t ldb, float *C, int64_t ldc, int F, float alpha, float beta,
                   const mggcn_epilogue *epi, const mggcn_spmm_plan_t *plan, void *ws) {
    const int nblk = plan ? (int)plan->blk.size() : 1;
    int w = F;
    if (nblk > 1) {
        const int64_t budget = resident_budget();
        w = 256;
        while (w > 32 && plan->blk_cols * (int64_t)w * 4 > budget) w /= 2;
        if (g_window > 0) w = g_window;
        if (w > F) w = F;
    } else if (g_slices > 1) {
        if (F % (4 * g_slices) == 0 && F / g_slices >= 32) w = F / g_slices;
    } else if (g_slices == 0) {
        const double bytes = (double)(t_touched_cols >= 0 ? t_touched_cols : cols) * (double)F * 4.0;
        const double l2 = (double)l2_bytes();
        if (bytes > 0.68 * l2 && bytes <= 0.94 * l2 && F % 8 == 0 && F / 2 >= 64) w = F / 2;
    }
    if (w % 4 != 0 && w != F) w = (w + 3) / 4 * 4;
    if (nblk == 1 && w >= F)
        return spmm_dispatch(st, rows, cols, rowptr, rowptr64, colidx, vals, B, ldb, C, ldc, F, alpha, beta, epi, plan, nullptr, ws);
    int nwin = 0;
    char first[sizeof t_last_config];
    first[0] = 0;  // F == 0 launches nothing
    for (int f0 = 0; f0 < F; f0 += w, ++nwin) {
        const int fw = std::min(w, F - f0);
        mggcn_epilogue e2{};
        if (epi) {
            e2 = *epi;
            if (e2.bias) e2.bias += f0;
            if (e2.aux) e2.aux += f0;
            if (e2.keep_mask) e2.keep_mask += f0;
        }
        for (int j = 0; j < nblk; ++j) {
            const bool last = j == nblk - 1;
            const int rc = spmm_dispatch(st, rows, cols, rowptr, rowptr64, colidx, vals, B + f0, ldb, C + f0, ldc, fw, alpha,
                                         j == 0 ? beta : 1.f, last && epi ? &e2 : nullptr, plan, plan ? &plan->blk[j] : nullptr, ws);
            if (rc) return rc;
            if (nwin == 0 && j == 0) memcpy(first, t_last_config, sizeof first);
        }
    }
    char tmp[sizeof t_last_config];
    snprintf(tmp, sizeof tmp, "%.150s x %d feature window(s) of %d x %d column block(s) of %lld", first, nwin, w, nblk,
             (long long)(plan && nblk > 1 ? plan->blk_cols : cols));
    memcpy(t_last_config, tmp, sizeof tmp);
    return MGGCN_OK;
}

int pick_split(uint64_t nnz, int split_nnz) {
    if (split_nnz <= 0) split_nnz = g_split;
    if (split_nnz <= 0) {
        // A warp works through its item with a fixed number of gathers in flight, so one long item
        // is a serial chain of L2 round trips: keep every item well below the kernel's duration.
        // Aim at >= 2 items per resident warp slot (148 SMs x 32 warps), between 64 and 2048 entries
        // (measured on B200: reddit-shape whole matrix best at 2048, its 8x8 blocks at 256-512).
        uint64_t s = nnz / (148ull * 32 * 2);
        s = (s + 31) / 32 * 32;
        split_nnz = (int)std::min<uint64_t>(2048, std::max<uint64_t>(64, s));
    }
    return std::max(split_nnz, 32);
}

// Appends one launch worth of items to `items` / `split`: rows r in [0, rows) with entries [beg(r), beg(r) + len(r));
// rows without entries are listed only when keep_empty.  Rows longer than split_nnz become equal segments (listed
// first), whole rows follow longest-first (counting sort on the length: stable, O(rows)).
template <typename beg_fn, typename len_fn>
plan_block build_block(int64_t rows, int split_nnz, bool keep_empty, beg_fn beg, len_fn len, std::vector<spmm_item> &items,
                       std::vector<split_row> &split) {
    plan_block pb;
    pb.first_item = (int64_t)items.size();
    pb.first_split = (int64_t)split.size();
    std::vector<uint32_t> whole;
    whole.reserve((size_t)rows);
    for (int64_t r = 0; r < rows; ++r) {
        const uint64_t l = len(r);
        pb.nnz += (int64_t)l;
        if (l > (uint64_t)split_nnz) {
            const uint64_t nseg = (l + split_nnz - 1) / split_nnz;
            const uint64_t seg = (l + nseg - 1) / nseg;
            split.push_back({(uint32_t)r, (uint32_t)(items.size() - (size_t)pb.first_item), (uint32_t)nseg});
            for (uint64_t k = 0; k < nseg; ++k) {
                const uint64_t b = k * seg, e = std::min(l, b + seg);
                items.push_back({beg(r) + b, (uint32_t)(e - b), (uint32_t)r});
            }
        } else if (l > 0 || keep_empty) {
            whole.push_back((uint32_t)r);
        }
    }
    pb.n_seg = (int64_t)items.size() - pb.first_item;
    pb.n_split = (int64_t)split.size() - pb.first_split;
    std::vector<uint64_t> hist((size_t)split_nnz + 2, 0);
    for (uint32_t r : whole) hist[(size_t)split_nnz - (size_t)len(r) + 1]++;
    for (size_t i = 1; i < hist.size(); ++i) hist[i] += hist[i - 1];
    const size_t base = items.size();
    items.resize(base + whole.size());
    for (uint32_t r : whole) {
        const uint64_t l = len(r);
        items[base + hist[(size_t)split_nnz - (size_t)l]++] = {beg(r), (uint32_t)l, r};
    }
    pb.n_items = (int64_t)items.size() - pb.first_item;
    return pb;
}

int plan_upload(mggcn_spmm_plan *p, const std::vector<spmm_item> &items, const std::vector<split_row> &split,
                mggcn_spmm_plan_t **out) {
    for (const plan_block &b : p->blk) p->max_seg = std::max(p->max_seg, b.n_seg);
    cudaError_t e = cudaGetDevice(&p->device);
    if (e == cudaSuccess && !items.empty()) e = cudaMalloc(&p->d_items, items.size() * sizeof(spmm_item));
    if (e == cudaSuccess && !items.empty())
        e = cudaMemcpy(p->d_items, items.data(), items.size() * sizeof(spmm_item), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && !split.empty()) e = cudaMalloc(&p->d_split, split.size() * sizeof(split_row));
    if (e == cudaSuccess && !split.empty())
        e = cudaMemcpy(p->d_split, split.data(), split.size() * sizeof(split_row), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        if (p->d_items) cudaFree(p->d_items);
        if (p->d_split) cudaFree(p->d_split);
        delete p;
        return fail((int)e, "spmm plan upload: %s", cudaGetErrorString(e));
    }
    *out = p;
    return MGGCN_OK;
}

template <typename idx_t>
int plan_create(int64_t rows, const idx_t *rowptr, int split_nnz, mggcn_spmm_plan_t **out) {
    if (!out || rows < 0 || (rows > 0 && !rowptr)) return fail(MGGCN_E_INVALID, "bad argument");
    if (rows > 0xFFFFFFFFll) return fail(MGGCN_E_UNSUPPORTED, "more than 2^32 rows in one block");
    split_nnz = pick_split(rows > 0 ? (uint64_t)rowptr[rows] - (uint64_t)rowptr[0] : 0, split_nnz);
    auto *p = new (std::nothrow) mggcn_spmm_plan;
    if (!p) return fail(MGGCN_E_NOMEM, "out of memory");
    p->rows = rows;
    p->split_nnz = split_nnz;
    std::vector<spmm_item> items;
    std::vector<split_row> split;
    p->blk.push_back(build_block(rows, split_nnz, true, [&](int64_t r) { return (uint64_t)rowptr[r]; },
                                 [&](int64_t r) { return (uint64_t)rowptr[r + 1] - (uint64_t)rowptr[r]; }, items, split));
    return plan_upload(p, items, split, out);
}

// DRAM bytes one product of width F is expected to move under a schedule of nblk column blocks (nblk = 1: none).
// Unblocked, a B larger than what stays resident misses on (1 - resident / B) of its gathers, each a sector read of
// the row slice; blocked, B is read once, the CSR stream once per feature window, and C once per block and window.
double schedule_cost(double rows, double cols, double nnz, int F, int nblk, double budget) {
    const double Fb = F * 4.0, Bbytes = cols * Fb;
    if (nblk <= 1) {
        const double miss = Bbytes <= budget ? 0.0 : 1.0 - budget / Bbytes;
        return nnz * 8.0 + Bbytes + nnz * Fb * miss + rows * Fb;
    }
    int w = 256;
    const double bc = std::ceil(cols / nblk);
    while (w > 32 && bc * w * 4.0 > budget) w /= 2;
    w = std::min(w, F);
    const double nwin = std::ceil((double)F / w);
    const double item_rows = rows * (2.0 + (nblk - 2) * std::min(1.0, nnz / rows / nblk));  // first and last list every row
    return nwin * (nnz * 8.0 + item_rows * 16.0) + Bbytes + rows * Fb * (2.0 * nblk - 1.0);
}

// Column-blocked plan.  Rows store their columns ascending in key(c) = (c - col_rot) mod cols (col_rot = 0: plain
// ascending CSR; the engine's grouped operands store stage blocks cyclically), so the entries of a row that fall
// into one column block are contiguous and a block is just another item list over the same colidx / vals arrays.
int plan_create_blocked(int64_t rows, int64_t cols, const uint32_t *rowptr, const uint32_t *colidx, uint32_t col_rot,
                        int64_t col_span, int F_hint, int split_nnz, mggcn_spmm_plan_t **out) {
    if (!out || rows < 0 || cols < 0 || (rows > 0 && !rowptr)) return fail(MGGCN_E_INVALID, "bad argument");
    if (cols == 0 || !colidx) return plan_create<uint32_t>(rows, rowptr, split_nnz, out);
    if (col_span <= 0 || col_span > cols) col_span = cols;
    const uint64_t nnz = rows > 0 ? (uint64_t)rowptr[rows] : 0;
    int nblk = g_blocks;
    if (nblk == 0) {
        // Measured on B200 (profiles/r02_spmm_products_variants.jsonl): the 2-D schedule does cut DRAM traffic
        // (products-shape, F = 256: ~70 GB instead of 125 GB per aggregation) but every configuration tried is
        // SLOWER (25-46 ms against 17.7 ms), because a row keeps only ~8 entries per column block and the launches
        // become latency-bound item loops (DRAM 25 % busy).  The planner therefore never blocks on its own; the
        // traffic model below is only consulted when mggcn_tune("blocks_model", 1) asks for it.
        nblk = 1;
    } else if (nblk < 0) {
        const double budget = (double)resident_budget();
        nblk = 1;
        double best = schedule_cost((double)rows, (double)col_span, (double)nnz, F_hint, 1, budget);
        for (int w = 32; w <= 128; w *= 2) {
            const int q = (int)std::ceil((double)col_span * w * 4.0 / budget);
            if (q < 2 || q > 64) continue;
            const double c = schedule_cost((double)rows, (double)col_span, (double)nnz, F_hint, q, budget);
            if (c < 0.8 * best) { best = c; nblk = q; }
        }
    }
    if (nblk <= 1 || rows == 0) return plan_create<uint32_t>(rows, rowptr, split_nnz, out);
    if (rows > 0xFFFFFFFFll) return fail(MGGCN_E_UNSUPPORTED, "more than 2^32 rows in one block");
    auto *p = new (std::nothrow) mggcn_spmm_plan;
    if (!p) return fail(MGGCN_E_NOMEM, "out of memory");
    p->rows = rows;
    p->blk_cols = (col_span + nblk - 1) / nblk;
    p->split_nnz = split_nnz = pick_split(nnz / (uint64_t)nblk, split_nnz);
    // ptr[r * (nblk + 1) + j] = first entry of row r whose key is >= j * blk_cols
    std::vector<uint32_t> ptr((size_t)rows * (size_t)(nblk + 1));
    const uint32_t ucols = (uint32_t)cols;
    int bad = 0;
#pragma omp parallel for schedule(static) reduction(max : bad)
    for (int64_t r = 0; r < rows; ++r) {
        uint32_t *pr = ptr.data() + (size_t)r * (size_t)(nblk + 1);
        const uint32_t end = rowptr[r + 1];
        uint64_t prev = 0;
        int j = 0;
        pr[0] = rowptr[r];
        for (uint32_t k = rowptr[r]; k < end; ++k) {
            const uint32_t c = colidx[k];
            const uint64_t key = c >= col_rot ? c - col_rot : (uint64_t)c + ucols - col_rot;
            if (key < prev || key >= (uint64_t)col_span) bad = 1;  // not ascending in the rotated order / outside the span
            prev = key;
            while (j + 1 < nblk && key >= (uint64_t)(j + 1) * (uint64_t)p->blk_cols) pr[++j] = k;
        }
        while (j < nblk) pr[++j] = end;
    }
    if (bad) {
        delete p;
        return fail(MGGCN_E_INVALID, "rows are not sorted by column (rotated by %u) or reach beyond the column span", col_rot);
    }
    std::vector<spmm_item> items;
    std::vector<split_row> split;
    for (int j = 0; j < nblk; ++j) {
        const size_t st = (size_t)(nblk + 1);
        p->blk.push_back(build_block(rows, split_nnz, j == 0 || j == nblk - 1,
                                     [&](int64_t r) { return (uint64_t)ptr[(size_t)r * st + j]; },
                                     [&](int64_t r) { return (uint64_t)(ptr[(size_t)r * st + j + 1] - ptr[(size_t)r * st + j]); }, items,
                                     split));
    }
    return plan_upload(p, items, split, out);
}

}  // namespace

namespace mggcn {
// engine.cu: rows of B the next SpMM call of this thread really gathers from (-1: all `cols` of them)
void spmm_touched_cols_hint(int64_t n) { t_touched_cols = n; }
// engine.cu: the staged exchange folded into one launch
int spmm_gathered(cudaStream_t st, int64_t rows, int64_t cols, const uint32_t *rowptr, const uint32_t *colidx,
                  const float *vals, float *gathered, int64_t ld, float *C, int64_t ldc, int F, float alpha,
                  float beta, const mggcn_epilogue *epi, const mggcn_spmm_plan_t *plan, void *ws, const spmm_gather &g) {
    if (plan && plan->blk.size() != 1) return fail(MGGCN_E_INVALID, "the gathered form takes an unblocked plan");
    return spmm_dispatch(st, rows, cols, rowptr, 0, colidx, vals, gathered, ld, C, ldc, F, alpha, beta, epi, plan, nullptr, ws, &g);
}
}  // namespace mggcn

extern "C" {

int mggcn_spmm_gathered_f32_i32(mggcn_stream_t stream, int64_t rows, int64_t cols, const uint32_t *rowptr,
                                const uint32_t *colidx, const float *vals, float *gathered, const float *B_own,
                                int64_t ld, float *C, int64_t ldc, int F, float alpha, float beta, int P, int rank,
                                const uint32_t *cuts, const uint32_t *stage_flags, int *err,
                                const mggcn_epilogue *epi, const mggcn_spmm_plan_t *plan, void *ws) {
    if (!cuts || P < 1 || P > 16 || rank < 0 || rank >= P) return fail(MGGCN_E_INVALID, "bad partition description");
    if (plan && plan->blk.size() != 1) return fail(MGGCN_E_INVALID, "the gathered form takes an unblocked plan");
    mggcn::spmm_gather g;
    g.flags = stage_flags;
    g.B_own = B_own;
    g.own_lo = cuts[rank];
    g.own_n = cuts[rank + 1] - cuts[rank];
    g.P = P;
    g.rank = rank;
    g.err = err;
    for (int i = 0; i < 16; ++i) g.cuts[i] = i < P ? cuts[i] : 0xFFFFFFFFu;
    return spmm_dispatch(static_cast<cudaStream_t>(stream), rows, cols, rowptr, 0, colidx, vals, gathered, ld, C, ldc, F,
                         alpha, beta, epi, plan, nullptr, ws, &g);
}

int mggcn_spmm_csr_f32_i32(mggcn_stream_t stream, int64_t rows, int64_t cols, const uint32_t *rowptr,
                           const uint32_t *colidx, const float *vals, const float *B, int64_t ldb,
                           float *C, int64_t ldc, int F, float alpha, float beta,
                           const mggcn_epilogue *epi, const mggcn_spmm_plan_t *plan, void *ws) {
    return spmm_scheduled(static_cast<cudaStream_t>(stream), rows, cols, rowptr, 0, colidx, vals, B, ldb,
                          C, ldc, F, alpha, beta, epi, plan, ws);
}

int mggcn_spmm_csr_f32_i64(mggcn_stream_t stream, int64_t rows, int64_t cols, const uint64_t *rowptr,
                           const uint32_t *colidx, const float *vals, const float *B, int64_t ldb,
                           float *C, int64_t ldc, int F, float alpha, float beta,
                           const mggcn_epilogue *epi, const mggcn_spmm_plan_t *plan, void *ws) {
    return spmm_scheduled(static_cast<cudaStream_t>(stream), rows, cols, rowptr, 1, colidx, vals, B, ldb,
                          C, ldc, F, alpha, beta, epi, plan, ws);
}

int mggcn_spmm_plan_create_i32(int64_t rows, const uint32_t *rowptr_host, int split_nnz,
                               mggcn_spmm_plan_t **out) {
    return plan_create<uint32_t>(rows, rowptr_host, split_nnz, out);
}
int mggcn_spmm_plan_create_i64(int64_t rows, const uint64_t *rowptr_host, int split_nnz,
                               mggcn_spmm_plan_t **out) {
    return plan_create<uint64_t>(rows, rowptr_host, split_nnz, out);
}
int mggcn_spmm_plan_create_blocked_i32(int64_t rows, int64_t cols, const uint32_t *rowptr_host,
                                       const uint32_t *colidx_host, uint32_t col_rot, int64_t col_span, int F_hint,
                                       int split_nnz, mggcn_spmm_plan_t **out) {
    return plan_create_blocked(rows, cols, rowptr_host, colidx_host, col_rot, col_span, F_hint, split_nnz, out);
}

size_t mggcn_spmm_plan_workspace_bytes(const mggcn_spmm_plan_t *plan, int F) {
    if (!plan || F <= 0) return 0;
    return (size_t)plan->max_seg * (size_t)((F + 3) / 4 * 4) * sizeof(float);
}
int mggcn_spmm_plan_blocks(const mggcn_spmm_plan_t *plan) { return plan ? (int)plan->blk.size() : 0; }

void mggcn_spmm_plan_destroy(mggcn_spmm_plan_t *plan) {
    if (!plan) return;
    if (plan->d_items) cudaFree(plan->d_items);
    if (plan->d_split) cudaFree(plan->d_split);
    delete plan;
}

const char *mggcn_last_spmm_config(void) { return t_last_config; }

int mggcn_tune(const char *key, int value) {
    if (!key) return MGGCN_E_INVALID;
    int *slot = nullptr;
    if (!strcmp(key, "unroll")) slot = &g_unroll;
    else if (!strcmp(key, "threads")) slot = &g_threads;
    else if (!strcmp(key, "policy")) slot = &g_policy;
    else if (!strcmp(key, "split")) slot = &g_split;
    else if (!strcmp(key, "slices")) slot = &g_slices;
    else if (!strcmp(key, "blocks")) slot = &g_blocks;
    else if (!strcmp(key, "blocks_model")) { const int old = g_blocks < 0; if (value >= 0) g_blocks = value ? -1 : 0; return old; }
    else if (!strcmp(key, "tma")) slot = &g_tma;
    else if (!strcmp(key, "narrow")) slot = &g_narrow;
    else if (!strcmp(key, "window")) slot = &g_window;
    else if (!strcmp(key, "tile_mb")) slot = &g_tile_mb;
    else if (!strcmp(key, "gemm_tc")) slot = &mggcn::g_gemm_tc;
    else if (!strcmp(key, "gemm_kgroup")) slot = &mggcn::g_gemm_kgroup;
    else if (!strcmp(key, "gemm_split")) slot = &mggcn::g_gemm_split;
    else if (!strcmp(key, "gemm_presplit")) slot = &mggcn::g_gemm_presplit;
    else if (!strcmp(key, "gemm_ts")) slot = &mggcn::g_gemm_ts;
    else if (!strcmp(key, "gemm_stages")) slot = &mggcn::g_gemm_stages;
    else if (!strcmp(key, "gemm_bn")) slot = &mggcn::g_gemm_bn;
    else if (!strcmp(key, "skip_exchange")) slot = &mggcn::g_skip_exchange;
    else if (!strcmp(key, "xstreams")) slot = &mggcn::g_xstreams;
    else if (!strcmp(key, "xpull")) slot = &mggcn::g_xpull;
    else if (!strcmp(key, "xgroups")) slot = &mggcn::g_xgroups;
    else if (!strcmp(key, "h2d_sm")) slot = &mggcn::g_h2d_sm;
    else if (!strcmp(key, "h2d_chunk_kb")) slot = &mggcn::g_h2d_chunk_kb;
    else if (!strcmp(key, "input_flip")) slot = &mggcn::g_input_flip;
    if (!slot) return fail(MGGCN_E_INVALID, "unknown tuning key %s", key);
    const int old = *slot;
    if (slot == &g_threads && (value < 32 || value > 256 || value % 32)) return fail(MGGCN_E_INVALID, "threads must be 32..256");
    if (value >= 0) *slot = value;
    return old;
}

}  // extern "C"
