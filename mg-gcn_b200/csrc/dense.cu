// Dense kernels around the aggregation: the H*W transform (FP32 SIMT form with fused bias /
// activation epilogue and a deterministic split-K path for G_W = X^T G), column sums, the
// element-wise kernels the reference launches separately, the fused softmax-cross-entropy loss and
// the fused Adam step.  Reference anchors are cited per entry point in include/mggcn.h.
//
// The tensor-core (tcgen05 / TMEM / TMA, 3xTF32) form of the transform lives in gemm_tc.cu and is
// taken whenever the operands are 16-byte aligned; this file keeps the FP32 SIMT form for
// unaligned / tiny products (the reference's cublasSgemm is FP32 SIMT too, src/cuda_utils.hpp:168-171).
#include <algorithm>

#include "common.hpp"

using namespace mggcn;

namespace mggcn {
// gemm_tc.cu: tcgen05 / TMEM / TMA form (3xTF32), used whenever the operands are 16-byte aligned
bool gemm_tc_eligible(int64_t M, int64_t N, int64_t K, const float *A, int64_t lda, const float *B, int64_t ldb);
size_t gemm_tc_workspace_bytes(int ta, int tb, int64_t M, int64_t N, int64_t K);
int gemm_tc(cudaStream_t st, int ta, int tb, int64_t M, int64_t N, int64_t K, float alpha, const float *A, int64_t lda,
            const float *B, int64_t ldb, float beta, float *C, int64_t ldc, const mggcn_epilogue &epi, void *ws,
            size_t ws_bytes);
int launch_splitk_reduce(cudaStream_t st, const float *part, int nz, int64_t M, int64_t N, float *C, int64_t ldc,
                         float alpha, float beta, const mggcn_epilogue &epi);
}  // namespace mggcn

namespace {

__device__ __forceinline__ float epilogue_f(float x, const mggcn_epilogue &e, int64_t row, int64_t col) {
    if (e.bias) x += __ldg(e.bias + col);
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

// ------------------------------------------------------------------------------------------------
// SIMT GEMM, row-major C[M x N] = op(A)[M x K] * op(B)[K x N].
//   A_KC: A is stored M x K (k contiguous);  else stored K x M (m contiguous, the "ta" case).
//   B_KC: B is stored N x K (k contiguous, the "tb" case);  else stored K x N (n contiguous).
// 256 threads, BM x BN x 16 tiles, 8 x (BN/16) outputs per thread, register-prefetched smem tiles.
// gridDim.z > 1: split-K, slice z writes its raw partial product to part + z*M*N (no epilogue).
// ------------------------------------------------------------------------------------------------
struct gemm_args {
    const float *A, *B;
    float *C;
    float *part;
    int64_t M, N, K, lda, ldb, ldc;
    int64_t kchunk;
    float alpha, beta;
    mggcn_epilogue epi;
    int vecA, vecB, vecC;  // 128-bit access allowed on that operand
};

constexpr int BK = 16;

template <int BMN, bool KC>
struct tile_loader {
    // KC:  tile is BMN rows x 16 k, k contiguous in memory  -> thread reads float4 along k, stores transposed
    // !KC: tile is 16 k x BMN, mn contiguous                -> thread reads float4 along mn, stores direct
    static constexpr int NV = BMN * BK / 4 / 256;  // float4 per thread
    float4 r[NV];
    __device__ __forceinline__ void load(const float *P, int64_t ld, int64_t mn0, int64_t mn_end, int64_t k0,
                                         int64_t k_end, bool vec) {
        const int t = threadIdx.x;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int idx = t + i * 256;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (KC) {
                const int rr = (idx & 31) + 32 * (idx / 128);
                const int kq = (idx >> 5) & 3;
                const int64_t mn = mn0 + rr, k = k0 + 4 * kq;
                if (mn < mn_end) {
                    const float *p = P + mn * ld + k;
                    if (vec && k + 3 < k_end) {
                        v = __ldg(reinterpret_cast<const float4 *>(p));
                    } else {
                        if (k + 0 < k_end) v.x = __ldg(p + 0);
                        if (k + 1 < k_end) v.y = __ldg(p + 1);
                        if (k + 2 < k_end) v.z = __ldg(p + 2);
                        if (k + 3 < k_end) v.w = __ldg(p + 3);
                    }
                }
            } else {
                const int kk = idx / (BMN / 4), q = idx % (BMN / 4);
                const int64_t k = k0 + kk, mn = mn0 + 4 * q;
                if (k < k_end) {
                    const float *p = P + k * ld + mn;
                    if (vec && mn + 3 < mn_end) {
                        v = __ldg(reinterpret_cast<const float4 *>(p));
                    } else {
                        if (mn + 0 < mn_end) v.x = __ldg(p + 0);
                        if (mn + 1 < mn_end) v.y = __ldg(p + 1);
                        if (mn + 2 < mn_end) v.z = __ldg(p + 2);
                        if (mn + 3 < mn_end) v.w = __ldg(p + 3);
                    }
                }
            }
            r[i] = v;
        }
    }
    __device__ __forceinline__ void store(float (*S)[BMN + 4]) const {
        const int t = threadIdx.x;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int idx = t + i * 256;
            if (KC) {
                const int rr = (idx & 31) + 32 * (idx / 128);
                const int kq = (idx >> 5) & 3;
                S[4 * kq + 0][rr] = r[i].x;
                S[4 * kq + 1][rr] = r[i].y;
                S[4 * kq + 2][rr] = r[i].z;
                S[4 * kq + 3][rr] = r[i].w;
            } else {
                const int kk = idx / (BMN / 4), q = idx % (BMN / 4);
                *reinterpret_cast<float4 *>(&S[kk][4 * q]) = r[i];
            }
        }
    }
};

template <bool A_KC, bool B_KC, int BM, int BN>
__global__ void __launch_bounds__(256) gemm_kernel(const gemm_args g) {
    constexpr int TM = BM / 16, TN = BN / 16;  // outputs per thread: TM x TN, in 4-wide groups
    __shared__ __align__(16) float As[2][BK][BM + 4];
    __shared__ __align__(16) float Bs[2][BK][BN + 4];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int64_t m0 = (int64_t)blockIdx.y * BM, n0 = (int64_t)blockIdx.x * BN;
    const int64_t kb = (int64_t)blockIdx.z * g.kchunk;
    const int64_t ke = min(g.K, kb + g.kchunk);

    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

    tile_loader<BM, A_KC> la;
    tile_loader<BN, B_KC> lb;
    if (kb < ke) {
        la.load(g.A, g.lda, m0, g.M, kb, ke, g.vecA);
        lb.load(g.B, g.ldb, n0, g.N, kb, ke, g.vecB);
        la.store(As[0]);
        lb.store(Bs[0]);
    }
    __syncthreads();
    int buf = 0;
    for (int64_t k0 = kb; k0 < ke; k0 += BK) {
        const bool more = k0 + BK < ke;
        if (more) {
            la.load(g.A, g.lda, m0, g.M, k0 + BK, ke, g.vecA);
            lb.load(g.B, g.ldb, n0, g.N, k0 + BK, ke, g.vecB);
        }
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            float a[TM], b[TN];
#pragma unroll
            for (int i = 0; i < TM / 4; ++i) {
                const float4 v = *reinterpret_cast<const float4 *>(&As[buf][k][ty * 4 + i * 64]);
                a[4 * i + 0] = v.x; a[4 * i + 1] = v.y; a[4 * i + 2] = v.z; a[4 * i + 3] = v.w;
            }
#pragma unroll
            for (int j = 0; j < TN / 4; ++j) {
                const float4 v = *reinterpret_cast<const float4 *>(&Bs[buf][k][tx * 4 + j * 64]);
                b[4 * j + 0] = v.x; b[4 * j + 1] = v.y; b[4 * j + 2] = v.z; b[4 * j + 3] = v.w;
            }
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        if (more) {
            la.store(As[buf ^ 1]);
            lb.store(Bs[buf ^ 1]);
        }
        __syncthreads();
        buf ^= 1;
    }

    const bool split = gridDim.z > 1;
    float *out = split ? g.part + (int64_t)blockIdx.z * g.M * g.N : g.C;
    const int64_t ldo = split ? g.N : g.ldc;
    const bool vec = split ? (g.N % 4 == 0) : (g.vecC != 0);
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const int64_t m = m0 + ty * 4 + (i / 4) * 64 + (i % 4);
        if (m >= g.M) continue;
#pragma unroll
        for (int j4 = 0; j4 < TN / 4; ++j4) {
            const int64_t n = n0 + tx * 4 + j4 * 64;
            if (n >= g.N) continue;
            float *c = out + m * ldo + n;
            float x[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) x[e] = acc[i][4 * j4 + e];
            if (split) {
                if (vec && n + 3 < g.N) {
                    *reinterpret_cast<float4 *>(c) = make_float4(x[0], x[1], x[2], x[3]);
                } else {
                    for (int e = 0; e < 4 && n + e < g.N; ++e) c[e] = x[e];
                }
                continue;
            }
            if (vec && n + 3 < g.N) {
                float4 old = make_float4(0.f, 0.f, 0.f, 0.f);
                if (g.beta != 0.f) old = *reinterpret_cast<const float4 *>(c);
                x[0] = g.alpha * x[0] + (g.beta != 0.f ? g.beta * old.x : 0.f);
                x[1] = g.alpha * x[1] + (g.beta != 0.f ? g.beta * old.y : 0.f);
                x[2] = g.alpha * x[2] + (g.beta != 0.f ? g.beta * old.z : 0.f);
                x[3] = g.alpha * x[3] + (g.beta != 0.f ? g.beta * old.w : 0.f);
#pragma unroll
                for (int e = 0; e < 4; ++e) x[e] = epilogue_f(x[e], g.epi, m, n + e);
                *reinterpret_cast<float4 *>(c) = make_float4(x[0], x[1], x[2], x[3]);
            } else {
                for (int e = 0; e < 4 && n + e < g.N; ++e) {
                    float v = g.alpha * x[e];
                    if (g.beta != 0.f) v += g.beta * c[e];
                    c[e] = epilogue_f(v, g.epi, m, n + e);
                }
            }
        }
    }
}

// C = alpha * sum_z part[z] + beta * C, slices added in z order
__global__ void splitk_reduce_kernel(const float *part, int nz, int64_t M, int64_t N, float *C, int64_t ldc,
                                     float alpha, float beta, const mggcn_epilogue epi) {
    const int64_t total = M * N;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int z = 0; z < nz; ++z) s += part[(int64_t)z * total + i];
        const int64_t m = i / N, n = i - m * N;
        float *c = C + m * ldc + n;
        float v = alpha * s;
        if (beta != 0.f) v += beta * *c;
        *c = epilogue_f(v, epi, m, n);
    }
}

int splitk_slices(int ta, int64_t M, int64_t N, int64_t K) {
    if (!ta) return 1;
    const int64_t tiles = ((M + 127) / 128) * ((N + 127) / 128);
    if (K < 4096 || tiles >= 148) return 1;
    int64_t z = (148 * 2 + tiles - 1) / tiles;       // ~2 CTAs per SM
    z = std::min<int64_t>(z, (K + 1023) / 1024);      // at least 1024 k per slice
    return (int)std::max<int64_t>(1, z);
}

template <bool A_KC, bool B_KC>
int gemm_launch(cudaStream_t st, gemm_args &g, int nz) {
    const bool narrow = g.N <= 64;
    dim3 grid((unsigned)((g.N + (narrow ? 63 : 127)) / (narrow ? 64 : 128)), (unsigned)((g.M + 127) / 128),
              (unsigned)nz);
    if (grid.y > 65535u) {
        return fail(MGGCN_E_UNSUPPORTED, "GEMM M too large for one launch");
    }
    if (narrow) gemm_kernel<A_KC, B_KC, 128, 64><<<grid, 256, 0, st>>>(g);
    else gemm_kernel<A_KC, B_KC, 128, 128><<<grid, 256, 0, st>>>(g);
    MGGCN_LAUNCH_CHECK();
    return MGGCN_OK;
}

// ------------------------------------------------------------------------------------------------
// G_b = 1^T G (src/gcn.hpp:288-296 does it with a GEMV against a vector of ones).  Two deterministic levels:
// a CTA of 32 x 8 threads owns rows_per_cta rows; thread (x, y) adds rows y, y+8, ... of a 128-bit column group
// (four independent accumulator sets keep four loads in flight), the 8 row groups are added in y order;
// stage 2 adds the per-CTA rows of `part` the same way (8 interleaved partial sums, then y order).
template <bool VEC4>
__global__ void __launch_bounds__(256) colsum_stage1(const float *G, int64_t ldg, int64_t rows, int64_t N,
                                                     int64_t rows_per_cta, float *part) {
    __shared__ float red[8][32][4 + 1];
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_cta;
    const int64_t r1 = min(rows, r0 + rows_per_cta);
    for (int64_t c0 = 0; c0 < N; c0 += 128) {
        const int64_t col = c0 + 4 * tx;
        float s[4][4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int e = 0; e < 4; ++e) s[u][e] = 0.f;
        if (col < N) {
            int64_t r = r0 + ty;
            if (VEC4 && col + 3 < N) {
                for (; r + 24 < r1; r += 32) {
                    float4 v[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) v[u] = *reinterpret_cast<const float4 *>(G + (r + 8 * u) * ldg + col);
#pragma unroll
                    for (int u = 0; u < 4; ++u) { s[u][0] += v[u].x; s[u][1] += v[u].y; s[u][2] += v[u].z; s[u][3] += v[u].w; }
                }
                for (; r < r1; r += 8) {
                    const float4 v = *reinterpret_cast<const float4 *>(G + r * ldg + col);
                    s[0][0] += v.x; s[0][1] += v.y; s[0][2] += v.z; s[0][3] += v.w;
                }
            } else {
                for (; r < r1; r += 8)
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                        if (col + e < N) s[0][e] += G[r * ldg + col + e];
            }
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) red[ty][tx][e] = (s[0][e] + s[1][e]) + (s[2][e] + s[3][e]);
        __syncthreads();
        if (ty == 0 && col < N) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                if (col + e >= N) break;
                float t = 0.f;
#pragma unroll
                for (int y = 0; y < 8; ++y) t += red[y][tx][e];
                part[(int64_t)blockIdx.x * N + col + e] = t;
            }
        }
        __syncthreads();
    }
}
__global__ void __launch_bounds__(256) colsum_stage2(const float *part, int nparts, int64_t N, float *out) {
    __shared__ float red[8][33];
    const int64_t col = (int64_t)blockIdx.x * 32 + threadIdx.x;
    float s = 0.f;
    if (col < N)
        for (int p = threadIdx.y; p < nparts; p += 8) s += part[(int64_t)p * N + col];
    red[threadIdx.y][threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.y == 0 && col < N) {
        float t = 0.f;
#pragma unroll
        for (int y = 0; y < 8; ++y) t += red[y][threadIdx.x];
        out[col] = t;
    }
}
int colsum_parts(int64_t rows) { return (int)std::max<int64_t>(1, std::min<int64_t>(148 * 4, (rows + 255) / 256)); }

// ------------------------------------------------------------------------------------------------
__global__ void leaky_fwd_kernel(const float *in, float *out, size_t n, float slope) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float x = in[i], ax = slope * x;
        out[i] = x > ax ? x : ax;
    }
}
__global__ void leaky_bwd_kernel(const float *in, const float *gi, float *go, size_t n, float slope) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        go[i] = in[i] > 0.f ? gi[i] : slope * gi[i];
}
__global__ void broadcast_rows_kernel(const float *row, float *mat, size_t rows, size_t m, int discard) {
    const size_t n = rows * m;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float b = row[i % m];
        mat[i] = discard ? b : mat[i] + b;
    }
}
__global__ void axpy_kernel(size_t n, float alpha, const float *x, float *y) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        y[i] = fmaf(alpha, x[i], y[i]);
}
// y = alpha * x + beta * y   (axpby of src/cuda_utils.cu:81-87: the SGD step and Adam's first moment)
__global__ void axpby_kernel(size_t n, float alpha, const float *x, float beta, float *y) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        y[i] = alpha * x[i] + beta * y[i];
}
unsigned ew_grid(size_t n) { return (unsigned)std::max<size_t>(1, std::min<size_t>((n + 255) / 256, 148 * 16)); }

// ------------------------------------------------------------------------------------------------
// Fused loss: one warp per row.  Same operation order as the reference chain
// max_rows -> subtract_rows_exp -> row sum -> scale_rows -> max_row_indices -> index_log_rows ->
// add_indexed_rows(-1) -> scal(1/N) -> is_equal  (src/gcn.hpp:738-766,890-930) with the fast-math
// intrinsics the reference build uses (CMakeLists.txt:57).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) softmax_xent_kernel(float *logits, int64_t ld, int64_t rows, int C,
                                                           const int32_t *labels, float inv_n, double *part) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int warps = blockDim.x >> 5;
    double loss = 0.0, correct = 0.0;
    for (int64_t r = (int64_t)blockIdx.x * warps + warp; r < rows; r += (int64_t)gridDim.x * warps) {
        float *h = logits + r * ld;
        float mx = -INFINITY;
        for (int c = lane; c < C; c += 32) mx = fmaxf(mx, h[c]);
#pragma unroll
        for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        float s = 0.f;
        for (int c = lane; c < C; c += 32) s += __expf(h[c] - mx);
#pragma unroll
        for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        const int y = labels[r];
        float best = -INFINITY, py = 0.f;
        int arg = 0x7fffffff;
        for (int c = lane; c < C; c += 32) {
            const float p = __fdividef(__expf(h[c] - mx), s);
            if (best < p) { best = p; arg = c; }  // first maximum, strict <  (cuda_utils.cu:119-133)
            if (c == y) py = p;
            h[c] = (p - (c == y ? 1.f : 0.f)) * inv_n;
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
            if (ob > best || (ob == best && oa < arg)) { best = ob; arg = oa; }
            py += __shfl_xor_sync(0xffffffffu, py, o);
        }
        if (lane == 0) {
            loss += (double)fabsf(__logf(py));
            correct += arg == y ? 1.0 : 0.0;
        }
    }
    __shared__ double sl[8], sc[8];
    if (lane == 0) { sl[warp] = loss; sc[warp] = correct; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < warps; ++w) { a += sl[w]; b += sc[w]; }
        part[2 * blockIdx.x + 0] = a;
        part[2 * blockIdx.x + 1] = b;
    }
}
__global__ void xent_reduce_kernel(const double *part, int n, double *out) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int i = 0; i < n; ++i) { a += part[2 * i]; b += part[2 * i + 1]; }
        out[0] = a;
        out[1] = b;
    }
}
int xent_ctas(int64_t rows) { return (int)std::max<int64_t>(1, std::min<int64_t>(148 * 4, (rows + 7) / 8)); }

// Fused Adam: Saxpy(wd) + axpby + aaxpby + adam_final of src/gcn.hpp:253-279 in one pass.
__global__ void adam_kernel(size_t n, float *p, float *g, float *m, float *v, float step_size, float beta1,
                            float beta2, float wd, float eps, float bc2, size_t n_decay) {
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
        const float denom = sqrtf(__fdividef(vi, bc2)) + eps;
        p[i] -= __fdividef(step_size * mi, denom);
    }
}

}  // namespace

namespace mggcn {
int launch_splitk_reduce(cudaStream_t st, const float *part, int nz, int64_t M, int64_t N, float *C, int64_t ldc,
                         float alpha, float beta, const mggcn_epilogue &epi) {
    splitk_reduce_kernel<<<ew_grid((size_t)(M * N)), 256, 0, st>>>(part, nz, M, N, C, ldc, alpha, beta, epi);
    MGGCN_LAUNCH_CHECK();
    return MGGCN_OK;
}
}  // namespace mggcn

// ------------------------------------------------------------------------------------------------
// Built-in scratch for mggcn_colsum_static_f32 (declared in include/mggcn.h, no reference counterpart).
// The array is part of the library image: every GPU that loads libmggcn.so carries these
// COLSUM_STATIC_MAX_PARTS x COLSUM_STATIC_MAX_N floats (2,424,832 bytes) until the process exits, used or not.
// In host code `g_colsum_scratch` is only a handle for the runtime's symbol table, never a pointer a kernel could
// follow -- its device address differs per GPU and is asked for once per device (colsum_static_scratch below).
// ------------------------------------------------------------------------------------------------
constexpr int64_t COLSUM_STATIC_MAX_N = 1024;        // widest G the entry point accepts
constexpr int COLSUM_STATIC_MAX_PARTS = 148 * 4;     // the most stage-1 CTAs colsum_parts() ever asks for
constexpr int COLSUM_STATIC_MAX_DEVICES = 64;
static __device__ __align__(256) float g_colsum_scratch[(size_t)COLSUM_STATIC_MAX_PARTS * COLSUM_STATIC_MAX_N];

namespace {
std::atomic<float *> g_colsum_scratch_addr[COLSUM_STATIC_MAX_DEVICES];

// Device address of g_colsum_scratch on the CURRENT device.  The first call per device asks the runtime
// (cudaGetSymbolAddress, which also makes sure the module is loaded there); later calls read the cached pointer.
// Two threads racing through the first call both get the same answer, so the cache needs no lock.
int colsum_static_scratch(float **out) {
    int dev = -1;
    MGGCN_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= COLSUM_STATIC_MAX_DEVICES)
        return fail(MGGCN_E_UNSUPPORTED, "colsum_static: device ordinal %d is beyond the %d this library keeps an address for",
                    dev, COLSUM_STATIC_MAX_DEVICES);
    float *p = g_colsum_scratch_addr[dev].load(std::memory_order_acquire);
    if (!p) {
        MGGCN_CUDA_TRY(cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_colsum_scratch));
        if (!p) return fail(MGGCN_E_STATE, "colsum_static: the runtime returned no address for the scratch array");
        g_colsum_scratch_addr[dev].store(p, std::memory_order_release);
    }
    *out = p;
    return MGGCN_OK;
}
}  // namespace

extern "C" {

size_t mggcn_gemm_workspace_bytes(int ta, int tb, int64_t M, int64_t N, int64_t K) {
    (void)tb;
    const int nz = splitk_slices(ta, M, N, K);
    const size_t simt = nz > 1 ? (size_t)nz * (size_t)M * (size_t)N * sizeof(float) : 0;
    return std::max(simt, gemm_tc_workspace_bytes(ta, tb, M, N, K));
}

int mggcn_gemm_f32(mggcn_stream_t stream, int ta, int tb, int64_t M, int64_t N, int64_t K, float alpha,
                   const float *A, int64_t lda, const float *B, int64_t ldb, float beta, float *C,
                   int64_t ldc, const mggcn_epilogue *epi, void *ws, size_t ws_bytes) {
    if (M < 0 || N < 0 || K < 0) return fail(MGGCN_E_INVALID, "negative dimension");
    if (M == 0 || N == 0) return MGGCN_OK;
    if (!A || !B || !C) return fail(MGGCN_E_INVALID, "null operand");
    if (lda < (ta ? M : K) || ldb < (tb ? K : N) || ldc < N) return fail(MGGCN_E_INVALID, "leading dimension too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    gemm_args g{};
    g.A = A; g.B = B; g.C = C;
    g.M = M; g.N = N; g.K = K; g.lda = lda; g.ldb = ldb; g.ldc = ldc;
    g.alpha = alpha; g.beta = beta;
    if (epi) g.epi = *epi;
    if (g.epi.kind == MGGCN_EPI_LEAKY_RELU_BWD && !g.epi.aux) return fail(MGGCN_E_INVALID, "LEAKY_RELU_BWD epilogue needs aux");
    if (gemm_tc_eligible(M, N, K, A, lda, B, ldb))
        return gemm_tc(st, ta, tb, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, g.epi, ws, ws_bytes);
    g.vecA = lda % 4 == 0 && (reinterpret_cast<uintptr_t>(A) & 15) == 0;
    g.vecB = ldb % 4 == 0 && (reinterpret_cast<uintptr_t>(B) & 15) == 0;
    g.vecC = ldc % 4 == 0 && (reinterpret_cast<uintptr_t>(C) & 15) == 0;
    int nz = splitk_slices(ta, M, N, K);
    if (nz > 1 && (!ws || ws_bytes < (size_t)nz * M * N * sizeof(float))) nz = 1;  // no workspace: one slice
    g.part = static_cast<float *>(ws);
    g.kchunk = nz > 1 ? ((K + nz - 1) / nz + BK - 1) / BK * BK : std::max<int64_t>(K, 1);
    if (nz > 1) nz = (int)((K + g.kchunk - 1) / g.kchunk);
    int rc;
    if (!ta && !tb) rc = gemm_launch<true, false>(st, g, nz);
    else if (!ta && tb) rc = gemm_launch<true, true>(st, g, nz);
    else if (ta && !tb) rc = gemm_launch<false, false>(st, g, nz);
    else rc = gemm_launch<false, true>(st, g, nz);
    if (rc) return rc;
    if (nz > 1) {
        const int64_t total = M * N;
        splitk_reduce_kernel<<<ew_grid((size_t)total), 256, 0, st>>>(g.part, nz, M, N, C, ldc, alpha, beta, g.epi);
        MGGCN_LAUNCH_CHECK();
    }
    return MGGCN_OK;
}

size_t mggcn_colsum_workspace_bytes(int64_t rows, int64_t N) {
    return (size_t)colsum_parts(rows) * (size_t)std::max<int64_t>(N, 0) * sizeof(float);
}

int mggcn_colsum_f32(mggcn_stream_t stream, int64_t rows, int64_t N, const float *G, int64_t ldg, float *out,
                     void *ws, size_t ws_bytes) {
    if (rows < 0 || N < 0) return fail(MGGCN_E_INVALID, "negative dimension");
    if (N == 0) return MGGCN_OK;
    if (!out || (!G && rows > 0)) return fail(MGGCN_E_INVALID, "null operand");
    if (ws_bytes < mggcn_colsum_workspace_bytes(rows, N) || !ws) return fail(MGGCN_E_INVALID, "colsum workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int parts = colsum_parts(rows);
    const int64_t rpc = (rows + parts - 1) / parts;
    const bool vec4 = ldg % 4 == 0 && (reinterpret_cast<uintptr_t>(G) & 15) == 0;
    if (vec4) colsum_stage1<true><<<parts, dim3(32, 8), 0, st>>>(G, ldg, rows, N, std::max<int64_t>(rpc, 1), static_cast<float *>(ws));
    else colsum_stage1<false><<<parts, dim3(32, 8), 0, st>>>(G, ldg, rows, N, std::max<int64_t>(rpc, 1), static_cast<float *>(ws));
    MGGCN_LAUNCH_CHECK();
    colsum_stage2<<<(unsigned)((N + 31) / 32), dim3(32, 8), 0, st>>>(static_cast<const float *>(ws), parts, N, out);
    MGGCN_LAUNCH_CHECK();
    return MGGCN_OK;
}

int mggcn_leaky_relu_forward_f32(mggcn_stream_t stream, const float *in, float *out, size_t size, float slope) {
    if (!size) return MGGCN_OK;
    if (!in || !out) return fail(MGGCN_E_INVALID, "null operand");
    leaky_fwd_kernel<<<ew_grid(size), 256, 0, static_cast<cudaStream_t>(stream)>>>(in, out, size, slope);
    MGGCN_LAUNCH_CHECK();
    return MGGCN_OK;
}

int mggcn_leaky_relu_backward_f32(mggcn_stream_t stream, const float *in, const float *G_in, float *G_out,
                                  size_t size, float slope) {
    if (!size) return MGGCN_OK;
    if (!in || !G_in || !G_out) return fail(MGGCN_E_INVALID, "null operand");
    leaky_bwd_kernel<<<ew_grid(size), 256, 0, static_cast<cudaStream_t>(stream)>>>(in, G_in, G_out, size, slope);
    MGGCN_LAUNCH_CHECK();
    return MGGCN_OK;
}

int mggcn_broadcast_rows_f32(mggcn_stream_t stream, const float *row, float *mat, size_t rows, size_t m,
                             int discard) {
    if (!rows || !m) return MGGCN_OK;
    if (!row || !mat) return fail(MGGCN_E_INVALID, "null operand");
    broadcast_rows_kernel<<<ew_grid(rows * m), 256, 0, static_cast<cudaStream_t>(stream)>>>(row, mat, rows, m, discard);
    MGGCN_LAUNCH_CHECK();
    return MGGCN_OK;
}

int mggcn_axpy_f32(mggcn_stream_t stream, size_t n, float alpha, const float *x, float *y) {
    if (!n) return MGGCN_OK;
    if (!x || !y) return fail(MGGCN_E_INVALID, "null operand");
    axpy_kernel<<<ew_grid(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(n, alpha, x, y);
    MGGCN_LAUNCH_CHECK();
    return MGGCN_OK;
}

int mggcn_axpby_f32(mggcn_stream_t stream, size_t n, float alpha, const float *x, float beta, float *y) {
    if (!n) return MGGCN_OK;
    if (!x || !y) return fail(MGGCN_E_INVALID, "null operand");
    axpby_kernel<<<ew_grid(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(n, alpha, x, beta, y);
    MGGCN_LAUNCH_CHECK();
    return MGGCN_OK;
}

size_t mggcn_softmax_xent_workspace_bytes(int64_t rows) { return (size_t)xent_ctas(rows) * 2 * sizeof(double); }

int mggcn_softmax_xent_f32(mggcn_stream_t stream, float *logits, int64_t ld, int64_t rows, int C,
                           const int32_t *labels, float inv_n, double *out, void *ws, size_t ws_bytes) {
    if (rows < 0 || C <= 0 || ld < C) return fail(MGGCN_E_INVALID, "bad shape");
    if (!out) return fail(MGGCN_E_INVALID, "null output");
    if (!ws || ws_bytes < mggcn_softmax_xent_workspace_bytes(rows)) return fail(MGGCN_E_INVALID, "loss workspace too small");
    if (rows > 0 && (!logits || !labels)) return fail(MGGCN_E_INVALID, "null operand");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int ctas = xent_ctas(rows);
    softmax_xent_kernel<<<ctas, 256, 0, st>>>(logits, ld, rows, C, labels, inv_n, static_cast<double *>(ws));
    MGGCN_LAUNCH_CHECK();
    xent_reduce_kernel<<<1, 32, 0, st>>>(static_cast<const double *>(ws), ctas, out);
    MGGCN_LAUNCH_CHECK();
    return MGGCN_OK;
}

int mggcn_adam_f32(mggcn_stream_t stream, size_t n, float *p, float *g, float *m, float *v, float lr,
                   float beta1, float beta2, float wd, float eps, float bc1, float bc2, size_t n_decay) {
    if (!n) return MGGCN_OK;
    if (!p || !g || !m || !v) return fail(MGGCN_E_INVALID, "null operand");
    adam_kernel<<<ew_grid(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(n, p, g, m, v, lr / bc1, beta1, beta2, wd,
                                                                        eps, bc2, n_decay);
    MGGCN_LAUNCH_CHECK();
    return MGGCN_OK;
}

// Column sums without a workspace argument and without any allocation: the per-CTA partial rows go to the
// library's own g_colsum_scratch.  Same two kernels, same grid and same summation order as mggcn_colsum_f32, so the
// two entry points return the same bits.  Contract (repeated in include/mggcn.h):
//  - N <= 1024; wider inputs are refused with MGGCN_E_INVALID before anything is enqueued;
//  - the device that owns `stream` must be the current device of the calling thread;
//  - the first call on each device looks the array's address up and must therefore not sit inside a stream
//    capture; make one call per device during start-up;
//  - nothing in the array has to be zero at launch: stage 1 stores every partial that stage 2 reads, including the
//    rows of CTAs whose row range is empty, so what the previous call left behind is overwritten, never read;
//  - ONE array per GPU serves all callers: two calls that may run at the same time (different streams, or different
//    host threads of this process, on one device) overwrite each other's partials.  Order them.
int mggcn_colsum_static_f32(mggcn_stream_t stream, int64_t rows, int64_t N, const float *G, int64_t ldg, float *out) {
    if (rows < 0 || N < 0) return fail(MGGCN_E_INVALID, "negative dimension");
    if (N == 0) return MGGCN_OK;
    if (!out || (!G && rows > 0)) return fail(MGGCN_E_INVALID, "null operand");
    if (N > COLSUM_STATIC_MAX_N)
        return fail(MGGCN_E_INVALID, "colsum_static: N = %lld is wider than the built-in scratch serves (at most %lld columns); "
                    "use mggcn_colsum_f32 with a caller-owned workspace", (long long)N, (long long)COLSUM_STATIC_MAX_N);
    const int parts = colsum_parts(rows);
    if (parts > COLSUM_STATIC_MAX_PARTS) return fail(MGGCN_E_STATE, "colsum_static: %d partial rows exceed the scratch", parts);
    float *scratch = nullptr;
    if (int rc = colsum_static_scratch(&scratch)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t rpc = (rows + parts - 1) / parts;
    const bool vec4 = ldg % 4 == 0 && (reinterpret_cast<uintptr_t>(G) & 15) == 0;
    if (vec4) colsum_stage1<true><<<parts, dim3(32, 8), 0, st>>>(G, ldg, rows, N, std::max<int64_t>(rpc, 1), scratch);
    else colsum_stage1<false><<<parts, dim3(32, 8), 0, st>>>(G, ldg, rows, N, std::max<int64_t>(rpc, 1), scratch);
    MGGCN_LAUNCH_CHECK();
    colsum_stage2<<<(unsigned)((N + 31) / 32), dim3(32, 8), 0, st>>>(scratch, parts, N, out);
    MGGCN_LAUNCH_CHECK();
    return MGGCN_OK;
}

}  // extern "C"
