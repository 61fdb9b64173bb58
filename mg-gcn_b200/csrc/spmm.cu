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
#include <numeric>
#include <vector>

#include "common.hpp"

using namespace mggcn;

namespace mggcn {
extern int g_gemm_tc, g_gemm_kgroup;  // gemm_tc.cu
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

struct mggcn_spmm_plan {
    int64_t rows = 0;
    int64_t n_items = 0;  // segments + whole rows
    int64_t n_seg = 0;    // items [0, n_seg) are segments; item k writes partial slot k
    int64_t n_split_rows = 0;
    int split_nnz = 0;
    int device = 0;
    spmm_item *d_items = nullptr;
    split_row *d_split = nullptr;
};

namespace {

int g_unroll = 4;     // gathers in flight per lane (4 or 8)
int g_threads = 64;   // CTA size
int g_policy = 1;     // 1: L2 evict_last for B, evict_first for the CSR stream; 0: plain loads
int g_split = 0;      // row-split threshold of new plans; 0 = chosen from the matrix (see plan_create)
int g_slices = 0;     // k > 1: run as k launches over F/k-wide column windows of B and C; 0: k = 2 when B (cols x F floats)
                      // is larger than half of the L2 but its half is not (see spmm_sliced); 1: never

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
    if (a.g.flags) {  // gathered multi-GPU operand
        if (exact) spmm_kernel<LANES, NCH, VEC, U4, true, true><<<gr, threads, 0, st>>>(a);
        else spmm_kernel<LANES, NCH, VEC, U4, false, true><<<gr, threads, 0, st>>>(a);
    } else if (g_unroll >= 8 && LANES >= 8) {
        if (exact) spmm_kernel<LANES, NCH, VEC, 8, true><<<gr, threads, 0, st>>>(a);
        else spmm_kernel<LANES, NCH, VEC, 8, false><<<gr, threads, 0, st>>>(a);
    } else {
        if (exact) spmm_kernel<LANES, NCH, VEC, U4, true><<<gr, threads, 0, st>>>(a);
        else spmm_kernel<LANES, NCH, VEC, U4, false><<<gr, threads, 0, st>>>(a);
    }
    MGGCN_LAUNCH_CHECK();
    return MGGCN_OK;
}

int spmm_dispatch(cudaStream_t st, int64_t rows, int64_t cols, const void *rowptr, int rowptr64,
                  const uint32_t *colidx, const float *vals, const float *B, int64_t ldb, float *C,
                  int64_t ldc, int F, float alpha, float beta, const mggcn_epilogue *epi,
                  const mggcn_spmm_plan_t *plan, void *ws, const mggcn::spmm_gather *gather = nullptr) {
    if (rows < 0 || cols < 0 || F < 0) return fail(MGGCN_E_INVALID, "negative dimension");
    if (rows == 0 || F == 0) return MGGCN_OK;
    if (!rowptr && !plan) return fail(MGGCN_E_INVALID, "rowptr is null");
    if (!C || (!B && cols > 0)) return fail(MGGCN_E_INVALID, "null operand");
    if (ldb < F || ldc < F) return fail(MGGCN_E_INVALID, "leading dimension smaller than F");
    if (rows > 0xFFFFFFFFll) return fail(MGGCN_E_UNSUPPORTED, "more than 2^32 rows in one block");
    if (plan && plan->rows != rows) return fail(MGGCN_E_INVALID, "plan was built for %lld rows, got %lld",
                                               (long long)plan->rows, (long long)rows);
    spmm_args a{};
    a.items = plan ? plan->d_items : nullptr;
    a.n_items = plan ? plan->n_items : rows;
    a.n_seg = plan ? plan->n_seg : 0;
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
    if (aligned) {
        if (F4 <= 2) rc = launch_cfg<2, 1, 4>(st, a);
        else if (F4 <= 4) rc = launch_cfg<4, 1, 4>(st, a);
        else if (F4 <= 8) rc = launch_cfg<8, 1, 4>(st, a);
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
    if (plan && plan->n_split_rows > 0) {
        const int64_t total = plan->n_split_rows * (int64_t)F;
        const int grid = (int)std::min<int64_t>((total + 255) / 256, 148 * 8);
        spmm_fixup_kernel<<<grid, 256, 0, st>>>(plan->d_split, plan->n_split_rows, a.partial, a.ldp, C,
                                                ldc, F, alpha, beta, a.epi);
        MGGCN_LAUNCH_CHECK();
    }
    return MGGCN_OK;
}

thread_local int64_t t_touched_cols = -1;

// Feature slicing: k launches over F/k-wide column windows (every window a multiple of 4 floats, >= 32 wide).
int spmm_sliced(cudaStream_t st, int64_t rows, int64_t cols, const void *rowptr, int rowptr64, const uint32_t *colidx,
                const float *vals, const float *B, int64_t ldb, float *C, int64_t ldc, int F, float alpha, float beta,
                const mggcn_epilogue *epi, const mggcn_spmm_plan_t *plan, void *ws) {
    // Why: B is the operand with reuse, and the two L2 halves of the chip each keep their own copy of a line, so
    // B stays resident only up to about half of the 126 MB.  At reddit shape (119 MB at F = 128) the unsliced
    // kernel re-reads B from DRAM 9 times over (10.7 GB per launch); it is not DRAM-bound, but inside a training
    // loop the board runs at its power cap and the DRAM traffic costs SM clock: two 64-wide windows measured
    // 21.3 ms/epoch at 1.78 GHz against 23.2 ms at 1.60 GHz.  The sums per output element are the same either way.
    // At 2 GPUs (blocks of 60 MB) slicing costs 2.5 % instead; the crossover between 60 and 119 MB has not been
    // measured, so the rule only fires from 90 MB (midpoint) up to the L2 size.  Callers whose CSR
    // touches only part of B's rows (the engine's grouped launches) say so through spmm_touched_cols_hint.
    int k = g_slices;
    if (k == 0) {
        const double bytes = (double)(t_touched_cols >= 0 ? t_touched_cols : cols) * (double)F * 4.0;
        k = (bytes > 90e6 && bytes <= 124e6 && F % 8 == 0 && F / 2 >= 64) ? 2 : 1;
    }
    if (k <= 1 || F % (4 * k) != 0 || F / k < 32)
        return spmm_dispatch(st, rows, cols, rowptr, rowptr64, colidx, vals, B, ldb, C, ldc, F, alpha, beta, epi, plan, ws);
    const int w = F / k;
    for (int i = 0; i < k; ++i) {
        mggcn_epilogue e2{};
        if (epi) {
            e2 = *epi;
            if (e2.bias) e2.bias += i * w;
            if (e2.aux) e2.aux += i * w;
            if (e2.keep_mask) e2.keep_mask += i * w;
        }
        const int rc = spmm_dispatch(st, rows, cols, rowptr, rowptr64, colidx, vals, B + i * w, ldb, C + i * w, ldc, w, alpha,
                                     beta, epi ? &e2 : nullptr, plan, ws);
        if (rc) return rc;
    }
    return MGGCN_OK;
}

template <typename idx_t>
int plan_create(int64_t rows, const idx_t *rowptr, int split_nnz, mggcn_spmm_plan_t **out) {
    if (!out || rows < 0 || (rows > 0 && !rowptr)) return fail(MGGCN_E_INVALID, "bad argument");
    if (rows > 0xFFFFFFFFll) return fail(MGGCN_E_UNSUPPORTED, "more than 2^32 rows in one block");
    if (split_nnz <= 0) split_nnz = g_split;
    if (split_nnz <= 0) {
        // A warp works through its item with a fixed number of gathers in flight, so one long item
        // is a serial chain of L2 round trips: keep every item well below the kernel's duration.
        // Aim at >= 2 items per resident warp slot (148 SMs x 32 warps), between 64 and 2048 entries
        // (measured on B200: reddit-shape whole matrix best at 2048, its 8x8 blocks at 256-512).
        const uint64_t nnz = rows > 0 ? (uint64_t)rowptr[rows] - (uint64_t)rowptr[0] : 0;
        uint64_t s = nnz / (148ull * 32 * 2);
        s = (s + 31) / 32 * 32;
        split_nnz = (int)std::min<uint64_t>(2048, std::max<uint64_t>(64, s));
    }
    if (split_nnz < 32) split_nnz = 32;
    auto *p = new (std::nothrow) mggcn_spmm_plan;
    if (!p) return fail(MGGCN_E_NOMEM, "out of memory");
    p->rows = rows;
    p->split_nnz = split_nnz;

    // whole rows, longest first (counting sort on the length: stable, O(rows))
    std::vector<uint32_t> order;
    std::vector<spmm_item> items;
    std::vector<split_row> split;
    std::vector<uint32_t> whole;
    whole.reserve((size_t)rows);
    for (int64_t r = 0; r < rows; ++r) {
        const uint64_t len = (uint64_t)rowptr[r + 1] - (uint64_t)rowptr[r];
        if (len > (uint64_t)split_nnz) {
            const uint64_t nseg = (len + split_nnz - 1) / split_nnz;
            const uint64_t seg = (len + nseg - 1) / nseg;
            split.push_back({(uint32_t)r, (uint32_t)items.size(), (uint32_t)nseg});
            for (uint64_t k = 0; k < nseg; ++k) {
                const uint64_t b = k * seg, e = std::min(len, b + seg);
                items.push_back({(uint64_t)rowptr[r] + b, (uint32_t)(e - b), (uint32_t)r});
            }
        } else {
            whole.push_back((uint32_t)r);
        }
    }
    p->n_seg = (int64_t)items.size();
    p->n_split_rows = (int64_t)split.size();
    {
        std::vector<uint64_t> hist((size_t)split_nnz + 2, 0);
        for (uint32_t r : whole) hist[(size_t)split_nnz - (size_t)(rowptr[r + 1] - rowptr[r]) + 1]++;
        for (size_t i = 1; i < hist.size(); ++i) hist[i] += hist[i - 1];
        const size_t base = items.size();
        items.resize(base + whole.size());
        for (uint32_t r : whole) {
            const uint64_t len = (uint64_t)rowptr[r + 1] - (uint64_t)rowptr[r];
            items[base + hist[(size_t)split_nnz - (size_t)len]++] = {(uint64_t)rowptr[r], (uint32_t)len, r};
        }
    }
    p->n_items = (int64_t)items.size();
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

}  // namespace

namespace mggcn {
// engine.cu: rows of B the next SpMM call of this thread really gathers from (-1: all `cols` of them)
void spmm_touched_cols_hint(int64_t n) { t_touched_cols = n; }
// engine.cu: the staged exchange folded into one launch
int spmm_gathered(cudaStream_t st, int64_t rows, int64_t cols, const uint32_t *rowptr, const uint32_t *colidx,
                  const float *vals, float *gathered, int64_t ld, float *C, int64_t ldc, int F, float alpha,
                  float beta, const mggcn_epilogue *epi, const mggcn_spmm_plan_t *plan, void *ws, const spmm_gather &g) {
    return spmm_dispatch(st, rows, cols, rowptr, 0, colidx, vals, gathered, ld, C, ldc, F, alpha, beta, epi, plan, ws, &g);
}
}  // namespace mggcn

extern "C" {

int mggcn_spmm_gathered_f32_i32(mggcn_stream_t stream, int64_t rows, int64_t cols, const uint32_t *rowptr,
                                const uint32_t *colidx, const float *vals, float *gathered, const float *B_own,
                                int64_t ld, float *C, int64_t ldc, int F, float alpha, float beta, int P, int rank,
                                const uint32_t *cuts, const uint32_t *stage_flags, int *err,
                                const mggcn_epilogue *epi, const mggcn_spmm_plan_t *plan, void *ws) {
    if (!cuts || P < 1 || P > 16 || rank < 0 || rank >= P) return fail(MGGCN_E_INVALID, "bad partition description");
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
                         alpha, beta, epi, plan, ws, &g);
}

int mggcn_spmm_csr_f32_i32(mggcn_stream_t stream, int64_t rows, int64_t cols, const uint32_t *rowptr,
                           const uint32_t *colidx, const float *vals, const float *B, int64_t ldb,
                           float *C, int64_t ldc, int F, float alpha, float beta,
                           const mggcn_epilogue *epi, const mggcn_spmm_plan_t *plan, void *ws) {
    return spmm_sliced(static_cast<cudaStream_t>(stream), rows, cols, rowptr, 0, colidx, vals, B, ldb,
                       C, ldc, F, alpha, beta, epi, plan, ws);
}

int mggcn_spmm_csr_f32_i64(mggcn_stream_t stream, int64_t rows, int64_t cols, const uint64_t *rowptr,
                           const uint32_t *colidx, const float *vals, const float *B, int64_t ldb,
                           float *C, int64_t ldc, int F, float alpha, float beta,
                           const mggcn_epilogue *epi, const mggcn_spmm_plan_t *plan, void *ws) {
    return spmm_sliced(static_cast<cudaStream_t>(stream), rows, cols, rowptr, 1, colidx, vals, B, ldb,
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

size_t mggcn_spmm_plan_workspace_bytes(const mggcn_spmm_plan_t *plan, int F) {
    if (!plan || F <= 0) return 0;
    return (size_t)plan->n_seg * (size_t)((F + 3) / 4 * 4) * sizeof(float);
}

void mggcn_spmm_plan_destroy(mggcn_spmm_plan_t *plan) {
    if (!plan) return;
    if (plan->d_items) cudaFree(plan->d_items);
    if (plan->d_split) cudaFree(plan->d_split);
    delete plan;
}

int mggcn_tune(const char *key, int value) {
    if (!key) return MGGCN_E_INVALID;
    int *slot = nullptr;
    if (!strcmp(key, "unroll")) slot = &g_unroll;
    else if (!strcmp(key, "threads")) slot = &g_threads;
    else if (!strcmp(key, "policy")) slot = &g_policy;
    else if (!strcmp(key, "split")) slot = &g_split;
    else if (!strcmp(key, "slices")) slot = &g_slices;
    else if (!strcmp(key, "gemm_tc")) slot = &mggcn::g_gemm_tc;
    else if (!strcmp(key, "gemm_kgroup")) slot = &mggcn::g_gemm_kgroup;
    if (!slot) return fail(MGGCN_E_INVALID, "unknown tuning key %s", key);
    const int old = *slot;
    if (slot == &g_threads && (value < 32 || value > 256 || value % 32)) return fail(MGGCN_E_INVALID, "threads must be 32..256");
    if (value >= 0) *slot = value;
    return old;
}

}  // extern "C"
