// Tensor-core form of the dense transform for sm_100a: tcgen05.mma (kind::tf32) with TMEM
// accumulators, TMA-fed shared-memory operand ring, 3xTF32 error compensation.
//
// Replaces cublasSgemm (FP32 SIMT, no TF32 math mode: src/cuda_utils.hpp:158-172) for the
// H*W (+b, activation), G*W^T (+activation gradient) and X^T*G products of src/gcn.hpp:191-296.
// The reference's result is an FP32-accurate product, and the parity bar is 1e-5 relative, so a
// plain TF32 product (10-bit mantissa, ~1e-3) is not admissible.  Every operand x is split into
//     hi = tf32(x),  lo = tf32(x - hi)
// and the product is accumulated as  lo_a*hi_b + hi_a*lo_b + hi_a*hi_b  (the dropped lo*lo term is
// 2^-22 relative).  Tensor-core accumulation chains are kept short: after `kgroup` k-blocks the
// TMEM accumulator is drained into FP32 registers (round-to-nearest adds) by the epilogue warps,
// so the rounding behaviour of the tensor-core adder never builds up over a long K (K is the number
// of vertices in X^T*G).
//
// Structure (one persistent CTA per SM, 448 threads):
//   warp 0      TMA producer: raw fp32 tiles of A and B -> smem (128-byte swizzle), mbarrier tx
//   warps 2-5   splitters: raw tile -> hi (in place) + lo (sibling buffer), generic->async fence
//   warp 1      MMA issuer (one lane): 3 x tcgen05.mma.kind::tf32 per 8-wide k step,
//               tcgen05.commit frees the smem stage / publishes the TMEM accumulator
//   warps 6-13  accumulate + epilogue: tcgen05.ld the 128 x BN accumulator (two warps per TMEM lane
//               quarter, half of the columns each), add into registers; at the end of a tile alpha/beta,
//               bias, LeakyReLU (forward or gradient), keep-mask, transposed through smem so that
//               global accesses are float4 pieces of whole rows
// Operand layouts: both the k-contiguous ("K-major") and the row-contiguous ("MN-major") form of
// each operand are fed straight from the caller's row-major arrays -- no transposed copies.
// Split-K (X^T*G): slice z of K writes its raw partial tile; the slices are added in z order by
// splitk_reduce (deterministic, no atomics).
#include <cuda.h>  // CUtensorMap and its enums; the encoder is resolved at run time (no -lcuda)

#include <algorithm>
#include <mutex>

#include "common.hpp"

using namespace mggcn;

namespace mggcn {
int g_gemm_tc = 1;       // 0: never use this path (mggcn_tune "gemm_tc")
int g_gemm_kgroup = 2;   // k-blocks per tensor-core accumulation chain (mggcn_tune "gemm_kgroup")
int g_gemm_split = 1;    // 0: hi = rna_tf32(x), lo = rna_tf32(x - hi);  1 (default): hi = x with the low 13 mantissa bits
                         //    cleared, lo = x - hi left as it is (the tensor core drops its low bits).  The two cvt.rna
                         //    per element run on the XU pipe (16 lanes per SM), which the splitter warps saturate; AND +
                         //    FADD do not: 0.978 -> 0.909 ms over the eight reddit-shape products, max error 3.2e-7 ->
                         //    5.5e-7 of the result's largest entry (profiles/r02_gemm_split_modes.jsonl)
                         //    The dropped terms are one-sided, but each is a fixed small fraction of its own signed
                         //    product, so they cancel with the sum: G_W = H^T G at K = 233,000 with sum|ab| = 111 x |sum ab|
                         //    is within 2.0e-6 (truncating) / 1.5e-6 (rna) of fp64 per entry, mean signed error 7e-9 of the
                         //    rms (tests/test_gpu_dense.py::test_gemm_weight_gradient_under_cancellation,
                         //    profiles/r03_gemm_split_cancellation.log) -- so the large-K path keeps the default too
int g_gemm_presplit = 1; // a small B operand (the weights of H*W and G*W^T) is split once per call into hi / lo arrays
                         //    in the workspace and both halves arrive by TMA: the splitter warps only handle A
int g_gemm_stages = 0;   // experiments: ring slots in use (0 = all), tile width override (0 = by N)
int g_gemm_bn = 0;
int g_gemm_ts = 1;       // A's hi / lo halves in tensor memory (TS form, see cfg); 0 = both operands in shared memory
}  // namespace mggcn

namespace {

constexpr int BM = 128;          // tile rows (UMMA M, all 128 TMEM lanes)
constexpr int BK = 32;           // fp32 per k-block = one 128-byte swizzle row
constexpr int UK = 8;            // tf32 UMMA K
constexpr int NTHREADS = 448;
constexpr int SPLIT_WARP0 = 2, EPI_WARP0 = 6;
constexpr uint32_t A_TILE_BYTES = BM * BK * 4;  // 16 KB
constexpr int STG_LD = 16;       // epilogue staging row stride (floats); 16-byte chunk c of row r sits at chunk c ^ ((r >> 1) & 3):
                                 // conflict-free for the row-per-lane writes and for the 8-rows-x-4-chunks reads

// TS: A's hi / lo halves live in tensor memory (tcgen05.mma with a TMEM A operand), B's in shared memory.
// The shared-memory data pipe is what bounds the all-smem (SS) form: ncu on the SS kernel shows the LSU side
// (splitter loads / stores, epilogue staging) at 59 % and the tensor-core operand reads at 36 % of the pipe's peak,
// 94 % together (profiles/r02_ncu_gemm_tc_n608.txt).  TS removes the hi / lo stores of A and half of the tensor
// core's operand reads; per k-block the pipe moves 64 KB instead of 144 KB.
template <int BN, bool TS>
struct cfg {
    static constexpr int STAGES = TS ? (BN == 128 ? 4 : 6) : (BN == 128 ? 3 : 4);
    static constexpr uint32_t B_TILE_BYTES = BN * BK * 4;
    static constexpr uint32_t A_BYTES = (TS ? 1 : 2) * A_TILE_BYTES;  // raw tile only / hi + lo
    static constexpr uint32_t STAGE_BYTES = A_BYTES + 2 * B_TILE_BYTES;
    static constexpr uint32_t STG_BYTES = 8 * 32 * STG_LD * 4;
    static constexpr uint32_t BAR_BYTES = 256;
    static constexpr uint32_t SMEM_BYTES = STAGES * STAGE_BYTES + STG_BYTES + BAR_BYTES + 1024;
    static constexpr int A_COL0 = 2 * BN;                          // TS: first TMEM column of the A ring
    static constexpr int A_COLS = 2 * BK;                          // TS: hi (32 columns) + lo (32 columns) per stage
    static constexpr int TMEM_COLS = TS ? 512 : 2 * BN;            // accumulators (+ A ring): a power of two
    static_assert(!TS || 2 * BN + STAGES * 2 * BK <= 512, "tensor memory has 512 columns");
    static_assert(8 * (3 * STAGES + 4) + 4 <= (int)BAR_BYTES, "barrier table");
};

struct tc_args {
    int64_t M, N, K;
    float *C;
    int64_t ldc;
    float *part;       // split-K partials: nz x M x N
    int nz;
    int64_t kchunk;    // multiple of BK
    int mt, nt;        // tiles along M, N
    int kgroup;
    int stages;        // > 0: use only this many slots of the shared-memory ring (mggcn_tune "gemm_stages")
    int split_mode;    // g_gemm_split
    int b_presplit;    // B's hi / lo halves sit behind each other in the workspace and arrive as one box of mapB
    int a_one, b_one;  // an MN-major operand whose extent is a multiple of 32 arrives as ONE 3-D box per k-block
                       // (chunk, k, 32 contiguous) instead of BM/32 (BN/32) 2-D boxes: the producer is one thread, and
                       // nine cp.async.bulk.tensor per k-block kept it busy 80 % of the time (ncu source page, round 2)
    float alpha, beta;
    mggcn_epilogue epi;
};

// ---------------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Bounded wait (about 2 s of SM clock): a protocol bug traps instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    long long t0 = 0;
    for (uint32_t spin = 0; !done; ++spin) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (!done && (spin & 1023u) == 1023u) {
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > 4000000000ll) {
                printf("mggcn gemm_tc: mbarrier wait timed out (block %d thread %d bar %u parity %u)\n", blockIdx.x,
                       threadIdx.x, bar, parity);
                __trap();
            }
        }
    }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
// one lane of a converged warp (the same lane for the same mask every time)
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tc_mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc,
                                               uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// 16 consecutive 32-bit columns of this thread's TMEM lane (32x32b: lane i of the warp <-> TMEM lane base + i)
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float *v) {
    const uint32_t *r = reinterpret_cast<const uint32_t *>(v);
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
          "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float *v) {
    uint32_t *r = reinterpret_cast<uint32_t *>(v);
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// shared-memory matrix descriptor, 128-byte swizzle (cute::UMMA::SmemDescriptor: start [0,14),
// leading byte offset [16,30), stride byte offset [32,46), version=1 [46,48), layout=2 [61,64))
// layout: 2 = SWIZZLE_128B (16-byte chunks, K-major operands), 1 = SWIZZLE_128B_BASE32B (32-byte chunks:
// the only shared-memory layout the tensor core accepts for an MN-major 32-bit operand)
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout) {
    uint64_t d = (uint64_t)((addr & 0x3FFFFu) >> 4);
    d |= (uint64_t)(lbo_bytes >> 4) << 16;
    d |= (uint64_t)(sbo_bytes >> 4) << 32;
    d |= 1ull << 46;
    d |= (uint64_t)layout << 61;
    return d;
}

// x = hi + lo with hi a TF32 number.  Default: both halves rounded to nearest (dropped terms 2^-22 relative).
// trunc: hi = x with the low 13 mantissa bits cleared, lo = x - hi exactly (<= 13 significant bits; the tensor core
// uses its top 11): dropped terms 2^-20 relative, a third of the instructions.
__device__ __forceinline__ void split_tf32(float x, float &h, float &l, bool trunc) {
    if (trunc) {
        h = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
        l = x - h;
    } else {
        uint32_t u;
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
        h = __uint_as_float(u);
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x - h));
        l = __uint_as_float(u);
    }
}

// B (rows x cols, pitch ldb) -> hi / lo arrays of pitch ld_out, pad columns zeroed
__global__ void presplit_kernel(const float *__restrict__ B, int64_t ldb, int64_t rows, int64_t cols, float *__restrict__ hi,
                                float *__restrict__ lo, int64_t ld_out, int trunc) {
    const int64_t total = rows * ld_out;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / ld_out, c = i - r * ld_out;
        float h = 0.f, l = 0.f;
        if (c < cols) split_tf32(B[r * ldb + c], h, l, trunc != 0);
        hi[i] = h;
        lo[i] = l;
    }
}

__device__ __forceinline__ float epi_apply(float x, const mggcn_epilogue &e, int64_t row, int64_t col, float bias) {
    x += bias;
    if (e.kind == MGGCN_EPI_LEAKY_RELU) {
        const float ax = e.slope * x;
        x = x > ax ? x : ax;
    } else if (e.kind == MGGCN_EPI_LEAKY_RELU_BWD) {
        const float z = __ldg(e.aux + row * e.ld_aux + col);
        x = z > 0.f ? x : e.slope * x;
    }
    if (e.keep_mask) x = e.keep_mask[row * e.ld_mask + col] ? x * e.keep_scale : 0.f;
    return x;
}

// ---------------------------------------------------------------------------------------------
// A_MN: A is stored K x M (m contiguous, the reference's A_T flag);  else M x K.
// B_MN: B is stored K x N (n contiguous, the plain case);            else N x K (the B_T flag).
// ---------------------------------------------------------------------------------------------
template <int BN, bool A_MN, bool B_MN, bool TS>
__global__ void __launch_bounds__(NTHREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, const tc_args g) {
    using C = cfg<BN, TS>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t *gen_base = smem_raw + (base - smem_u32(smem_raw));
    const uint32_t stg_addr = base + C::STAGES * C::STAGE_BYTES;
    float *stg = reinterpret_cast<float *>(gen_base + C::STAGES * C::STAGE_BYTES);
    const uint32_t bars = stg_addr + C::STG_BYTES;
    // barrier table (8 bytes each): full[S], conv[S], empty[S], tfull[2], tempty[2]; then the TMEM pointer
    auto bar_full = [&](int s) { return bars + 8u * s; };
    auto bar_conv = [&](int s) { return bars + 8u * (C::STAGES + s); };
    auto bar_empty = [&](int s) { return bars + 8u * (2 * C::STAGES + s); };
    auto bar_tfull = [&](int a) { return bars + 8u * (3 * C::STAGES + a); };
    auto bar_tempty = [&](int a) { return bars + 8u * (3 * C::STAGES + 2 + a); };
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(gen_base + C::STAGES * C::STAGE_BYTES +
                                                                         C::STG_BYTES + 8 * (3 * C::STAGES + 4));
    // stage layout: SS = [A hi (raw, split in place) | A lo | B hi | B lo];  TS = [A raw | B hi | B lo]
    auto a_hi = [&](int s) { return base + s * C::STAGE_BYTES; };
    auto a_lo = [&](int s) { return base + s * C::STAGE_BYTES + A_TILE_BYTES; };  // SS only
    auto b_hi = [&](int s) { return base + s * C::STAGE_BYTES + C::A_BYTES; };
    auto b_lo = [&](int s) { return base + s * C::STAGE_BYTES + C::A_BYTES + C::B_TILE_BYTES; };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nst = g.stages > 0 && g.stages < C::STAGES ? g.stages : C::STAGES;  // ring depth in use (experiments)

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&mapA)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&mapB)) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < C::STAGES; ++s) {
            mbar_init(bar_full(s), 1);
            mbar_init(bar_conv(s), 4);
            mbar_init(bar_empty(s), 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(bar_tfull(a), 1);
            mbar_init(bar_tempty(a), 8);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         smem_u32(const_cast<uint32_t *>(tmem_slot))),
                     "r"((uint32_t)C::TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int64_t tiles_mn = (int64_t)g.mt * g.nt;
    const int64_t total = tiles_mn * g.nz;
    const int64_t kb_all = (g.K + BK - 1) / BK;
    const int64_t kb_per_z = g.kchunk / BK;

    if (warp == 0) {
        // ================= TMA producer =================
        if (lane == 0) {
            int s = 0;
            uint32_t ph = 0;
            for (int64_t t = blockIdx.x; t < total; t += gridDim.x) {
                const int z = (int)(t / tiles_mn);
                const int64_t mn = t - (int64_t)z * tiles_mn;
                const int m0 = (int)(mn % g.mt) * BM, n0 = (int)(mn / g.mt) * BN;
                const int64_t kb0 = z * kb_per_z, kb1 = min(kb_all, kb0 + kb_per_z);
                for (int64_t kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(bar_empty(s), ph ^ 1u);
                    mbar_expect_tx(bar_full(s), A_TILE_BYTES + (g.b_presplit ? 2u : 1u) * C::B_TILE_BYTES);
                    const int k0 = (int)(kb * BK);
                    if (A_MN) {
                        if (g.a_one) {
                            tma_load_3d(a_hi(s), &mapA, bar_full(s), 0, k0, m0 / 32);
                        } else {
#pragma unroll
                            for (int j = 0; j < BM / 32; ++j) tma_load_2d(a_hi(s) + j * 4096, &mapA, bar_full(s), m0 + 32 * j, k0);
                        }
                    } else {
                        tma_load_2d(a_hi(s), &mapA, bar_full(s), k0, m0);
                    }
                    if (g.b_presplit) {  // hi and lo in one box: [half][chunk][k][32 n] or [half][n][k]
                        if (B_MN) tma_load_4d(b_hi(s), &mapB, bar_full(s), 0, k0, n0 / 32, 0);
                        else tma_load_3d(b_hi(s), &mapB, bar_full(s), k0, n0, 0);
                    } else if (B_MN) {
                        if (g.b_one) {
                            tma_load_3d(b_hi(s), &mapB, bar_full(s), 0, k0, n0 / 32);
                        } else {
#pragma unroll
                            for (int j = 0; j < BN / 32; ++j) tma_load_2d(b_hi(s) + j * 4096, &mapB, bar_full(s), n0 + 32 * j, k0);
                        }
                    } else {
                        tma_load_2d(b_hi(s), &mapB, bar_full(s), k0, n0);
                    }
                    if (++s == nst) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        // The WHOLE warp walks the loop with warp-uniform state and one elected lane issues: operands that the
        // compiler can prove uniform live in uniform registers.  Under `if (lane == 0)` every tcgen05.mma was wrapped
        // in an ELECT + 4 x R2UR.BROADCAST + branch sequence (per-thread registers -> uniform operands), about
        // 140 cycles per instruction against 60 of tensor-pipe time: the issuing thread, not the tensor core, not
        // shared memory and not the ring depth, bounded the kernel (profiles/r02_ncu_gemm_tc_ts_form.txt and the stage / tile-width sweeps of round 2).
        {
            const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);  // provably uniform copy
            // instruction descriptor (cute::UMMA::InstrDescriptor): D=f32 [4,6), A=tf32 [7,10), B=tf32 [10,13),
            // a_major [15], b_major [16], N>>3 [17,23), M>>4 [24,29)
            // (a TMEM A operand is always K-major: lane = row, column = k)
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((A_MN && !TS ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                                   ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
            // K-major (SWIZZLE_128B): 8-row groups 1024 B apart (SBO), k step = 32 B inside the swizzle row.
            // MN-major (SWIZZLE_128B_BASE32B, atom = 4 k-rows x 128 B): 32-column chunks 4096 B apart (LBO),
            // 4-k-row groups 512 B apart (SBO), k step of 8 rows = 1024 B.
            constexpr uint32_t A_LBO = A_MN ? 4096 : 16, A_SBO = A_MN ? 512 : 1024, A_STEP = A_MN ? 1024 : 32, A_LAY = A_MN ? 1 : 2;
            constexpr uint32_t B_LBO = B_MN ? 4096 : 16, B_SBO = B_MN ? 512 : 1024, B_STEP = B_MN ? 1024 : 32, B_LAY = B_MN ? 1 : 2;
            int s = 0, a = 0;
            uint32_t ph = 0, aph = 0;
            for (int64_t t = blockIdx.x; t < total; t += gridDim.x) {
                const int z = (int)(t / tiles_mn);
                const int64_t kb0 = z * kb_per_z, kb1 = min(kb_all, kb0 + kb_per_z);
                for (int64_t kg = kb0; kg < kb1; kg += g.kgroup) {
                    const int64_t kge = min(kb1, kg + g.kgroup);
                    mbar_wait(bar_tempty(a), aph ^ 1u);
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + (uint32_t)(a * BN);
                    for (int64_t kb = kg; kb < kge; ++kb) {
                        mbar_wait(bar_conv(s), ph);
                        tc_fence_after();
                        const uint64_t bh = smem_desc(b_hi(s), B_LBO, B_SBO, B_LAY), bl = smem_desc(b_lo(s), B_LBO, B_SBO, B_LAY);
                        if constexpr (TS) {
                            // A: TMEM columns [A_COL0 + 64 s, +32) = hi, the next 32 = lo; a k step is 8 columns
                            const uint32_t th = tmem_base + (uint32_t)(C::A_COL0 + s * C::A_COLS), tl = th + (uint32_t)BK;
#pragma unroll
                            for (int k = 0; k < BK / UK; ++k) {
                                const uint64_t bo = (uint64_t)((k * B_STEP) >> 4);
                                if (elect_one()) {
                                    tc_mma_tf32_ts(d_tmem, tl + (uint32_t)(k * UK), bh + bo, idesc, (kb > kg || k > 0) ? 1u : 0u);
                                    tc_mma_tf32_ts(d_tmem, th + (uint32_t)(k * UK), bl + bo, idesc, 1u);
                                    tc_mma_tf32_ts(d_tmem, th + (uint32_t)(k * UK), bh + bo, idesc, 1u);
                                }
                            }
                        } else {
                            const uint64_t ah = smem_desc(a_hi(s), A_LBO, A_SBO, A_LAY), al = smem_desc(a_lo(s), A_LBO, A_SBO, A_LAY);
#pragma unroll
                            for (int k = 0; k < BK / UK; ++k) {
                                const uint64_t ao = (uint64_t)((k * A_STEP) >> 4), bo = (uint64_t)((k * B_STEP) >> 4);
                                if (elect_one()) {
                                    tc_mma_tf32(d_tmem, al + ao, bh + bo, idesc, (kb > kg || k > 0) ? 1u : 0u);
                                    tc_mma_tf32(d_tmem, ah + ao, bl + bo, idesc, 1u);
                                    tc_mma_tf32(d_tmem, ah + ao, bh + bo, idesc, 1u);
                                }
                            }
                        }
                        // (the elected lane is the same one every time: commits track the MMAs that lane issued)
                        if (elect_one()) tc_commit(bar_empty(s));  // smem stage may be refilled once these MMAs have read it
                        __syncwarp();
                        if (++s == nst) { s = 0; ph ^= 1u; }
                    }
                    if (elect_one()) tc_commit(bar_tfull(a));      // accumulator chain complete
                    __syncwarp();
                    if (++a == 2) { a = 0; aph ^= 1u; }
                }
            }
        }
    } else if (warp < EPI_WARP0) {
        // ================= splitters: raw -> hi (in place), lo =================
        const int tid = threadIdx.x - SPLIT_WARP0 * 32;  // 0..127
        int s = 0;
        uint32_t ph = 0;
        for (int64_t t = blockIdx.x; t < total; t += gridDim.x) {
            const int z = (int)(t / tiles_mn);
            const int64_t kb0 = z * kb_per_z, kb1 = min(kb_all, kb0 + kb_per_z);
            for (int64_t kb = kb0; kb < kb1; ++kb) {
                mbar_wait(bar_full(s), ph);
                const bool trunc = g.split_mode != 0;
                auto split4 = [trunc](float4 *ph4, float4 *pl4) {
                    float4 x = *ph4, h, l;
                    split_tf32(x.x, h.x, l.x, trunc);
                    split_tf32(x.y, h.y, l.y, trunc);
                    split_tf32(x.z, h.z, l.z, trunc);
                    split_tf32(x.w, h.w, l.w, trunc);
                    *ph4 = h;
                    *pl4 = l;
                };
                if constexpr (TS) {
                    // thread <-> tile row <-> TMEM lane (a warp reaches the lane quarter warp % 4): read the row's 32
                    // raw values of this k-block from shared memory, write hi / lo into the stage's TMEM columns
                    tc_fence_after();
                    const int q = warp & 3, row = q * 32 + lane;
                    const uint8_t *araw = gen_base + s * C::STAGE_BYTES;
                    const uint32_t t_hi = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(C::A_COL0 + s * C::A_COLS);
                    float x[BK];
                    if constexpr (!A_MN) {
                        // K-major tile: row r = 128 bytes, 16-byte chunk c at position c ^ (r % 8) (TMA SWIZZLE_128B)
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float4 v = *reinterpret_cast<const float4 *>(araw + row * 128 + ((j ^ (row & 7)) << 4));
                            x[4 * j] = v.x; x[4 * j + 1] = v.y; x[4 * j + 2] = v.z; x[4 * j + 3] = v.w;
                        }
                    } else {
                        // MN-major tile: four unswizzled boxes of 32 k-rows x 32 m; this warp's rows are box q
#pragma unroll
                        for (int k = 0; k < BK; ++k) x[k] = *reinterpret_cast<const float *>(araw + q * 4096 + k * 128 + lane * 4);
                    }
#pragma unroll
                    for (int hf = 0; hf < 2; ++hf) {
                        float h[16], l[16];
#pragma unroll
                        for (int k = 0; k < 16; ++k) split_tf32(x[hf * 16 + k], h[k], l[k], trunc);
                        tmem_st16(t_hi + (uint32_t)(hf * 16), h);
                        tmem_st16(t_hi + (uint32_t)(BK + hf * 16), l);
                    }
                    tmem_st_wait();
                } else {
                    float4 *hi = reinterpret_cast<float4 *>(gen_base + s * C::STAGE_BYTES);
                    float4 *lo = reinterpret_cast<float4 *>(gen_base + s * C::STAGE_BYTES + A_TILE_BYTES);
#pragma unroll
                    for (int i = 0; i < (int)(A_TILE_BYTES / 16 / 128); ++i) split4(hi + tid + 128 * i, lo + tid + 128 * i);
                }
                if (!g.b_presplit) {
                    float4 *bhi = reinterpret_cast<float4 *>(gen_base + s * C::STAGE_BYTES + C::A_BYTES);
                    float4 *blo = reinterpret_cast<float4 *>(gen_base + s * C::STAGE_BYTES + C::A_BYTES + C::B_TILE_BYTES);
#pragma unroll
                    for (int i = 0; i < (int)(C::B_TILE_BYTES / 16 / 128); ++i) split4(bhi + tid + 128 * i, blo + tid + 128 * i);
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic writes -> tensor-core reads
                if constexpr (TS) tc_fence_before();                          // tcgen05.st -> the issuer's tcgen05.mma
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_conv(s));
                if (++s == nst) { s = 0; ph ^= 1u; }
            }
        }
    } else {
        // ================= accumulate + epilogue =================
        // 8 warps: warp w may read TMEM lanes 32*(w%4)..+31 only, so two warps share each lane quarter
        // and take half of the tile's columns each.
        constexpr int CW = BN / 2;           // columns per warp
        const int q = warp & 3;              // TMEM lane quarter (= 32 tile rows)
        const int ew = warp - EPI_WARP0;     // staging slot 0..7
        const int half = ew >> 2;
        float *my_stg = stg + ew * 32 * STG_LD;
        int a = 0;
        uint32_t aph = 0;
        for (int64_t t = blockIdx.x; t < total; t += gridDim.x) {
            const int z = (int)(t / tiles_mn);
            const int64_t mn = t - (int64_t)z * tiles_mn;
            const int64_t m0 = (int64_t)(mn % g.mt) * BM, n0 = (int64_t)(mn / g.mt) * BN + half * CW;
            const int64_t kb0 = z * kb_per_z, kb1 = min(kb_all, kb0 + kb_per_z);
            float acc[CW];
#pragma unroll
            for (int i = 0; i < CW; ++i) acc[i] = 0.f;
            for (int64_t kg = kb0; kg < kb1; kg += g.kgroup) {
                mbar_wait(bar_tfull(a), aph);
                tc_fence_after();
#pragma unroll
                for (int c = 0; c < CW / 32; ++c) {
                    float v[32];
                    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(a * BN + half * CW + c * 32), v);
#pragma unroll
                    for (int j = 0; j < 32; ++j) acc[c * 32 + j] += v[j];
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_tempty(a));
                if (++a == 2) { a = 0; aph ^= 1u; }
            }
            // store: rows m0 + 32q .. +31, 16 columns at a time through the staging slot.  The slot is
            // written thread-per-row (what tcgen05.ld delivers) and read back 8 rows x 4 float4 per
            // instruction: every global access of the warp is made of whole 64-byte row pieces and one lane
            // finishes 4 elements per instruction.  Operands of the epilogue (the activation input of the
            // LeakyReLU gradient, beta*C) are loaded for all row groups of a chunk before its first store.
            const bool split = g.nz > 1;
            float *out = split ? g.part + (int64_t)z * g.M * g.N : g.C;
            const int64_t ldo = split ? g.N : g.ldc;
            const bool use_aux = !split && g.epi.kind == MGGCN_EPI_LEAKY_RELU_BWD;
            const bool use_old = !split && g.beta != 0.f;
            const bool vec_ok = (reinterpret_cast<uintptr_t>(out) & 15) == 0 && ldo % 4 == 0 &&
                                (split || (!g.epi.keep_mask &&
                                           (!use_aux || ((reinterpret_cast<uintptr_t>(g.epi.aux) & 15) == 0 && g.epi.ld_aux % 4 == 0))));
            const int r_in = lane >> 2, c4 = (lane & 3) * 4;
            const int64_t row0 = m0 + q * 32 + r_in;
#pragma unroll
            for (int c = 0; c < CW / 16; ++c) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    *reinterpret_cast<float4 *>(my_stg + lane * STG_LD + 4 * (j ^ ((lane >> 1) & 3))) =
                        make_float4(acc[c * 16 + 4 * j], acc[c * 16 + 4 * j + 1], acc[c * 16 + 4 * j + 2], acc[c * 16 + 4 * j + 3]);
                __syncwarp();
                const int64_t col = n0 + c * 16 + c4;
                if (vec_ok && col + 3 < g.N) {
                    float4 bias = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (!split && g.epi.bias) bias = make_float4(__ldg(g.epi.bias + col), __ldg(g.epi.bias + col + 1),
                                                                 __ldg(g.epi.bias + col + 2), __ldg(g.epi.bias + col + 3));
                    float4 zz[4], old[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) zz[i] = old[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (use_aux) {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            if (row0 + 8 * i < g.M)
                                zz[i] = __ldg(reinterpret_cast<const float4 *>(g.epi.aux + (row0 + 8 * i) * g.epi.ld_aux + col));
                    }
                    if (use_old) {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            if (row0 + 8 * i < g.M) old[i] = *reinterpret_cast<const float4 *>(out + (row0 + 8 * i) * ldo + col);
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int64_t row = row0 + 8 * i;
                        if (row >= g.M) break;
                        float4 x4 = *reinterpret_cast<const float4 *>(my_stg + (8 * i + r_in) * STG_LD + (c4 ^ (((r_in >> 1) & 3) << 2)));
                        if (!split) {
                            float x[4] = {x4.x, x4.y, x4.z, x4.w};
                            const float o[4] = {old[i].x, old[i].y, old[i].z, old[i].w};
                            const float zv[4] = {zz[i].x, zz[i].y, zz[i].z, zz[i].w};
                            const float bv[4] = {bias.x, bias.y, bias.z, bias.w};
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                float y = x[e] * g.alpha;
                                if (use_old) y += g.beta * o[e];
                                y += bv[e];
                                if (g.epi.kind == MGGCN_EPI_LEAKY_RELU) {
                                    const float ay = g.epi.slope * y;
                                    y = y > ay ? y : ay;
                                } else if (use_aux) {
                                    y = zv[e] > 0.f ? y : g.epi.slope * y;
                                }
                                x[e] = y;
                            }
                            x4 = make_float4(x[0], x[1], x[2], x[3]);
                        }
                        *reinterpret_cast<float4 *>(out + row * ldo + col) = x4;
                    }
                } else if (col < g.N) {
                    for (int i = 0; i < 4; ++i) {
                        const int64_t row = row0 + 8 * i;
                        if (row >= g.M) break;
                        for (int e = 0; e < 4 && col + e < g.N; ++e) {
                            float x = my_stg[(8 * i + r_in) * STG_LD + (c4 ^ (((r_in >> 1) & 3) << 2)) + e];
                            float *cp = out + row * ldo + col + e;
                            if (!split) {
                                x *= g.alpha;
                                if (g.beta != 0.f) x += g.beta * *cp;
                                x = epi_apply(x, g.epi, row, col + e, g.epi.bias ? __ldg(g.epi.bias + col + e) : 0.f);
                            }
                            *cp = x;
                        }
                    }
                }
                __syncwarp();
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)C::TMEM_COLS)
                     : "memory");
    }
}

// ---------------------------------------------------------------------------------------------
// host
// ---------------------------------------------------------------------------------------------
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
encode_tiled_fn g_encode = nullptr;
int g_sm_count = 0;
bool g_tc_usable = false;
std::once_flag g_once;

void tc_init() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return;
    int major = 0;
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn) {
        cudaGetLastError();
        return;
    }
    g_encode = reinterpret_cast<encode_tiled_fn>(fn);
    g_tc_usable = major == 10 && g_sm_count > 0;
}

// 2-D fp32 tensor map over a row-major array: `inner` contiguous elements, `outer` rows of `ld` elements.
// mn_major: the operand's M/N index is the contiguous one -> 32-byte-atom 128-byte swizzle (see smem_desc)
// plain: no swizzle (a tile that only threads read: the MN-major A operand of the TS form)
int make_map(CUtensorMap *m, const float *p, int64_t inner, int64_t outer, int64_t ld, uint32_t box_inner,
             uint32_t box_outer, bool mn_major, bool plain = false) {
    cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    cuuint32_t box[2] = {box_inner, box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(p), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE,
                          plain ? CU_TENSOR_MAP_SWIZZLE_NONE : mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(MGGCN_E_INVALID, "cuTensorMapEncodeTiled failed (%d)", (int)r);
    return MGGCN_OK;
}

// n-dimensional view (rank 3 or 4) for operands that arrive as one box per k-block; strides in elements
int make_map_nd(CUtensorMap *m, const float *p, int rank, const int64_t *dims, const int64_t *strides, const uint32_t *box,
                CUtensorMapSwizzle sw) {
    cuuint64_t d[4], st[3];
    cuuint32_t b[4], es[4] = {1, 1, 1, 1};
    for (int i = 0; i < rank; ++i) { d[i] = (cuuint64_t)dims[i]; b[i] = box[i]; }
    for (int i = 0; i + 1 < rank; ++i) st[i] = (cuuint64_t)strides[i] * 4;
    CUresult r = g_encode(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<float *>(p), d, st, b, es,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(MGGCN_E_INVALID, "cuTensorMapEncodeTiled (rank %d) failed (%d)", rank, (int)r);
    return MGGCN_OK;
}

struct split_plan {
    int nz;
    int64_t kchunk;
};
split_plan plan_split(int ta, int64_t M, int64_t N, int64_t K, int bn) {
    const int64_t kb_all = (K + BK - 1) / BK;
    split_plan p{1, std::max<int64_t>(kb_all, 1) * BK};
    if (!ta || K < 4096) return p;
    const int64_t tiles = ((M + BM - 1) / BM) * ((N + bn - 1) / bn);
    const int sms = g_sm_count > 0 ? g_sm_count : 148;
    int64_t nz = std::max<int64_t>(1, sms / tiles);
    nz = std::min<int64_t>(nz, std::max<int64_t>(1, kb_all / 8));  // at least 8 k-blocks per slice
    if (nz <= 1) return p;
    const int64_t kb_per = (kb_all + nz - 1) / nz;
    p.kchunk = kb_per * BK;
    p.nz = (int)((kb_all + kb_per - 1) / kb_per);
    return p;
}

template <int BN, bool A_MN, bool B_MN, bool TS>
int launch(cudaStream_t st, const CUtensorMap &ma, const CUtensorMap &mb, const tc_args &g) {
    using C = cfg<BN, TS>;
    // the attribute is per device: one process may drive several GPUs (the mg_gcn CLI does)
    static std::atomic<bool> attr_set[64];
    int dev = 0;
    MGGCN_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev].load(std::memory_order_acquire)) {
        MGGCN_CUDA_TRY(cudaFuncSetAttribute(gemm_tc_kernel<BN, A_MN, B_MN, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)C::SMEM_BYTES));
        if (dev >= 0 && dev < 64) attr_set[dev].store(true, std::memory_order_release);
    }
    const int64_t total = (int64_t)g.mt * g.nt * g.nz;
    const unsigned grid = (unsigned)std::min<int64_t>(total, g_sm_count);
    gemm_tc_kernel<BN, A_MN, B_MN, TS><<<grid, NTHREADS, C::SMEM_BYTES, st>>>(ma, mb, g);
    MGGCN_LAUNCH_CHECK();
    return MGGCN_OK;
}

}  // namespace

namespace mggcn {

int launch_splitk_reduce(cudaStream_t st, const float *part, int nz, int64_t M, int64_t N, float *C, int64_t ldc,
                         float alpha, float beta, const mggcn_epilogue &epi);

bool gemm_tc_eligible(int64_t M, int64_t N, int64_t K, const float *A, int64_t lda, const float *B, int64_t ldb) {
    std::call_once(g_once, tc_init);
    if (!g_tc_usable || !g_gemm_tc) return false;
    if (M < 1 || N < 1 || K < 1 || N > 128 * 65535ll) return false;
    if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15) || lda % 4 || ldb % 4) return false;
    if (M * N * K < (1ll << 18)) return false;  // tiny products: the SIMT kernel has no setup cost
    return true;
}

// The weight operand is pre-split when the product is tall (its cost is amortised over many row tiles) and B is small
bool presplit_wanted(int ta, int64_t M, int64_t N, int64_t K) {
    return g_gemm_presplit && !ta && M >= 4096 && K * ((N + 31) / 32 * 32) <= (1ll << 21) && N * ((K + 3) / 4 * 4) <= (1ll << 21);
}
// pitch of the hi / lo copies: the MN-major form (K x N) is read in 32-column chunks, so its rows are padded to 32
int64_t presplit_ld(int tb, int64_t N, int64_t K) { return tb ? (K + 3) / 4 * 4 : (N + 31) / 32 * 32; }
size_t presplit_bytes(int tb, int64_t N, int64_t K) {
    const int64_t rows = tb ? N : K;
    return (size_t)(2 * rows * presplit_ld(tb, N, K)) * sizeof(float) + 256;
}

size_t gemm_tc_workspace_bytes(int ta, int tb, int64_t M, int64_t N, int64_t K) {
    std::call_once(g_once, tc_init);
    const split_plan p = plan_split(ta, M, N, K, N <= 64 ? 64 : 128);
    size_t need = p.nz > 1 ? (size_t)p.nz * (size_t)M * (size_t)N * sizeof(float) : 0;
    if (presplit_wanted(ta, M, N, K)) need = (need + 255) / 256 * 256 + presplit_bytes(tb, N, K);
    return need;
}

int gemm_tc(cudaStream_t st, int ta, int tb, int64_t M, int64_t N, int64_t K, float alpha, const float *A, int64_t lda,
            const float *B, int64_t ldb, float beta, float *C, int64_t ldc, const mggcn_epilogue &epi, void *ws,
            size_t ws_bytes) {
    const int bn = g_gemm_bn == 64 || g_gemm_bn == 128 ? g_gemm_bn : N <= 64 ? 64 : 128;
    split_plan sp = plan_split(ta, M, N, K, bn);
    if (sp.nz > 1 && (!ws || ws_bytes < (size_t)sp.nz * M * N * sizeof(float))) {
        sp.nz = 1;
        sp.kchunk = (K + BK - 1) / BK * BK;
    }
    // pre-split weight operand: hi / lo arrays behind the split-K partials (never both in practice)
    const size_t part_bytes = sp.nz > 1 ? ((size_t)sp.nz * M * N * sizeof(float) + 255) / 256 * 256 : 0;
    const float *Bs = B;
    int64_t ldbs = ldb;
    const bool presplit = presplit_wanted(ta, M, N, K) && ws && ws_bytes >= part_bytes + presplit_bytes(tb, N, K);
    const int64_t b_rows = tb ? N : K, b_cols = tb ? K : N;
    if (presplit) {
        ldbs = presplit_ld(tb, N, K);
        char *w = static_cast<char *>(ws) + part_bytes;
        w += (256 - (reinterpret_cast<uintptr_t>(w) & 255)) & 255;
        float *hi = reinterpret_cast<float *>(w), *lo = hi + b_rows * ldbs;
        const int64_t total = b_rows * ldbs;
        presplit_kernel<<<(unsigned)std::min<int64_t>((total + 255) / 256, 1024), 256, 0, st>>>(B, ldb, b_rows, b_cols, hi, lo, ldbs,
                                                                                             g_gemm_split);
        MGGCN_LAUNCH_CHECK();
        Bs = hi;
    }
    CUtensorMap ma, mb;
    const bool ts = g_gemm_ts != 0;
    // A: ta -> stored K x M (inner = M): one 3-D box (32 m, 32 k, 4 chunks) when M is a multiple of 32, else four
    // 2-D boxes of 32 m x 32 k; not ta -> stored M x K (inner = K), one box 32 k x 128 m
    const bool a_one = ta && M % 32 == 0;
    if (a_one) {
        const int64_t dims[3] = {32, K, M / 32}, str[2] = {lda, 32};
        const uint32_t box[3] = {32, (uint32_t)BK, (uint32_t)(BM / 32)};
        int rc = make_map_nd(&ma, A, 3, dims, str, box, ts ? CU_TENSOR_MAP_SWIZZLE_NONE : CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
        if (rc) return rc;
    } else if (ta) {
        int rc = make_map(&ma, A, M, K, lda, 32, BK, true, ts); if (rc) return rc;
    } else {
        int rc = make_map(&ma, A, K, M, lda, BK, BM, false); if (rc) return rc;
    }
    // B: tb -> stored N x K (inner = K), box 32 k x BN n; else stored K x N (inner = N), 32-column chunks.
    // The pre-split copies add an outermost dimension of 2 (hi, lo) so that both halves are one box.
    bool b_one = false;
    if (presplit && tb) {
        const int64_t dims[3] = {K, N, 2}, str[2] = {ldbs, b_rows * ldbs};
        const uint32_t box[3] = {(uint32_t)BK, (uint32_t)bn, 2};
        int rc = make_map_nd(&mb, Bs, 3, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B); if (rc) return rc;
    } else if (presplit) {
        const int64_t dims[4] = {32, K, ldbs / 32, 2}, str[3] = {ldbs, 32, b_rows * ldbs};
        const uint32_t box[4] = {32, (uint32_t)BK, (uint32_t)(bn / 32), 2};
        int rc = make_map_nd(&mb, Bs, 4, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B); if (rc) return rc;
    } else if (tb) {
        int rc = make_map(&mb, B, K, N, ldb, BK, (uint32_t)bn, false); if (rc) return rc;
    } else if (N % 32 == 0) {
        b_one = true;
        const int64_t dims[3] = {32, K, N / 32}, str[2] = {ldb, 32};
        const uint32_t box[3] = {32, (uint32_t)BK, (uint32_t)(bn / 32)};
        int rc = make_map_nd(&mb, B, 3, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B); if (rc) return rc;
    } else {
        int rc = make_map(&mb, B, N, K, ldb, 32, BK, true); if (rc) return rc;
    }
    tc_args g{};
    g.M = M; g.N = N; g.K = K;
    g.C = C; g.ldc = ldc;
    g.part = static_cast<float *>(ws);
    g.nz = sp.nz;
    g.kchunk = sp.kchunk;
    g.mt = (int)((M + BM - 1) / BM);
    g.nt = (int)((N + bn - 1) / bn);
    g.kgroup = std::max(1, g_gemm_kgroup);
    g.stages = g_gemm_stages;
    g.split_mode = g_gemm_split;
    g.b_presplit = presplit ? 1 : 0;
    g.a_one = a_one ? 1 : 0;
    g.b_one = b_one ? 1 : 0;
    g.alpha = alpha; g.beta = beta;
    g.epi = epi;
    int rc;
#define MGGCN_TC_LAUNCH(BN_, TS_)                                                            \
    (ta && !tb   ? launch<BN_, true, true, TS_>(st, ma, mb, g)                           \
     : ta && tb  ? launch<BN_, true, false, TS_>(st, ma, mb, g)                          \
     : !ta && tb ? launch<BN_, false, false, TS_>(st, ma, mb, g)                         \
                 : launch<BN_, false, true, TS_>(st, ma, mb, g))
    if (bn == 64) rc = ts ? MGGCN_TC_LAUNCH(64, true) : MGGCN_TC_LAUNCH(64, false);
    else rc = ts ? MGGCN_TC_LAUNCH(128, true) : MGGCN_TC_LAUNCH(128, false);
#undef MGGCN_TC_LAUNCH
    if (rc) return rc;
    if (sp.nz > 1) return launch_splitk_reduce(st, g.part, sp.nz, M, N, C, ldc, alpha, beta, epi);
    return MGGCN_OK;
}

}  // namespace mggcn
