// Synthetic dataset generator: stands in for /root/reference/test/data/prep.py (which downloads
// reddit / ogbn-* through dgl and ogb; no network here) and writes the same shapes:
//   prep.py:101-126  N and F padded up to multiples of 8 with isolated, zero-feature, label-0
//                    nodes; one self-loop per node (add_self_loop); unit weights;
//   prep.py:87-94    optional seeded symmetric permutation ("test/data/permuted/<name>").
// Edges are sampled with a counter-based hash, so the graph depends only on the arguments, never
// on the thread count.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <limits>
#include <omp.h>
#include <random>
#include <vector>

#include "common.hpp"

using namespace mggcn;

struct mggcn_gen {
    int64_t n = 0;
    std::vector<uint64_t> rowptr;
    std::vector<uint32_t> colidx;
};

namespace {

// new_id = perm[old_id]; identity when seed == 0 (prep.py: seed 0 means "not permuted")
std::vector<uint32_t> make_perm(int64_t n, uint64_t seed) {
    std::vector<uint32_t> perm((size_t)n);
    for (int64_t i = 0; i < n; ++i) perm[i] = (uint32_t)i;
    if (seed == 0) return perm;
    for (int64_t i = n - 1; i > 0; --i) {  // Fisher-Yates driven by the counter hash
        uint64_t j = hash2(seed ^ 0x5045524Dull, (uint64_t)i) % (uint64_t)(i + 1);
        std::swap(perm[i], perm[j]);
    }
    return perm;
}

// Vose alias table over n weights; probability and alias of a slot share one 8-byte entry, so a sample costs one
// cache miss (the table of a 111 M-vertex graph is 0.9 GB and every sample is a random access)
struct alias_table {
    struct slot_t { float prob; uint32_t alias; };
    std::vector<slot_t> slots;
    void build(const std::vector<double> &w) {
        const size_t n = w.size();
        slots.assign(n, slot_t{0.f, 0});
        double total = 0;
        for (double x : w) total += x;
        std::vector<double> scaled(n);
        std::vector<uint32_t> small, large;
        small.reserve(n);
        large.reserve(n);
        for (size_t i = 0; i < n; ++i) {
            scaled[i] = w[i] * (double)n / total;
            (scaled[i] < 1.0 ? small : large).push_back((uint32_t)i);
        }
        while (!small.empty() && !large.empty()) {
            uint32_t s = small.back(), l = large.back();
            small.pop_back();
            slots[s].prob = (float)scaled[s];
            slots[s].alias = l;
            scaled[l] = (scaled[l] + scaled[s]) - 1.0;
            if (scaled[l] < 1.0) { large.pop_back(); small.push_back(l); }
        }
        for (uint32_t i : large) { slots[i].prob = 1.f; slots[i].alias = i; }
        for (uint32_t i : small) { slots[i].prob = 1.f; slots[i].alias = i; }
    }
    inline uint32_t sample(uint64_t h) const {
        const size_t n = slots.size();
        uint64_t slot = (h >> 32) * (uint64_t)n >> 32;           // top 32 bits pick the slot
        float u = (float)(h & 0xFFFFFFull) * (1.0f / 16777216.0f);  // low 24 bits pick the side
        const slot_t s = slots[slot];
        return u < s.prob ? (uint32_t)slot : s.alias;
    }
};

}  // namespace

// degree_model 2 -- the union of k affine bijections of Z_n and their inverses: the neighbours of u are
// a_j*u + b_j and a_j^-1 * (u - b_j) (mod n_raw), j < k = floor(directed edges / 2n), plus the self-loop.  The
// pattern is symmetric by construction and a row costs O(k) whichever rank builds it: no walk over a global sample
// stream (3.2 G samples per rank at papers100M-shape took minutes), nothing to exchange.  Every vertex has degree
// <= 2k + 1 (uniform like model 0's expectation); the multipliers are ~n-sized, so consecutive rows gather rows of B
// that are far apart -- the same no-reuse access pattern as a random graph.
static int gen_rows_affine(int64_t n_raw, int64_t n, int64_t target_directed_edges, uint64_t seed, int permute, int64_t r0,
                           int64_t r1, mggcn_gen_t **out, int64_t *n_padded, int64_t *nnz) {
    if (permute) return fail(MGGCN_E_UNSUPPORTED, "degree model 2 is generated unpermuted (its ids are already scattered)");
    const int64_t nl = r1 - r0;
    const int k = (int)std::max<int64_t>(1, target_directed_edges / (2 * n_raw));
    const uint64_t m = (uint64_t)n_raw;
    std::vector<uint64_t> a((size_t)k), ainv((size_t)k), b((size_t)k);
    auto gcd = [](uint64_t x, uint64_t y) { while (y) { uint64_t t = x % y; x = y; y = t; } return x; };
    for (int j = 0; j < k; ++j) {
        uint64_t aj = m > 2 ? 2 + hash2(seed ^ 0x41464649ull, (uint64_t)j) % (m - 2) : 1;
        while (gcd(aj, m) != 1) aj = aj + 1 < m ? aj + 1 : 1;
        // inverse by the extended Euclidean algorithm (signed 128-bit: the coefficients stay below m)
        __int128 t0 = 0, t1 = 1, r0e = (__int128)m, r1e = (__int128)aj;
        while (r1e != 0) {
            __int128 q = r0e / r1e, t2 = t0 - q * t1, r2 = r0e - q * r1e;
            t0 = t1; t1 = t2; r0e = r1e; r1e = r2;
        }
        a[j] = aj;
        ainv[j] = (uint64_t)((t0 % (__int128)m + (__int128)m) % (__int128)m);
        b[j] = hash2(seed ^ 0x4F464653ull, (uint64_t)j) % m;
    }
    auto row = [&](int64_t u, uint32_t *buf) -> uint32_t {  // sorted, deduplicated, self-loop included
        if (u >= n_raw) { buf[0] = (uint32_t)u; return 1; }  // padding vertex: isolated but for its self-loop
        uint32_t c = 0;
        for (int j = 0; j < k; ++j) {
            const uint64_t v1 = (a[j] * (uint64_t)u + b[j]) % m;
            const uint64_t v2 = ainv[j] * (((uint64_t)u + m - b[j]) % m) % m;
            buf[c++] = (uint32_t)v1;
            buf[c++] = (uint32_t)v2;
        }
        buf[c++] = (uint32_t)u;
        std::sort(buf, buf + c);
        return (uint32_t)(std::unique(buf, buf + c) - buf);
    };
    auto *g = new (std::nothrow) mggcn_gen;
    if (!g) return fail(MGGCN_E_NOMEM, "out of memory");
    g->n = nl;
    g->rowptr.assign((size_t)nl + 1, 0);
#pragma omp parallel
    {
        std::vector<uint32_t> buf((size_t)2 * k + 1);
#pragma omp for schedule(static)
        for (int64_t i = 0; i < nl; ++i) g->rowptr[i + 1] = row(r0 + i, buf.data());
    }
    for (int64_t i = 0; i < nl; ++i) g->rowptr[i + 1] += g->rowptr[i];
    g->colidx.resize(g->rowptr[nl]);
#pragma omp parallel
    {
        std::vector<uint32_t> buf((size_t)2 * k + 1);
#pragma omp for schedule(static)
        for (int64_t i = 0; i < nl; ++i) {
            const uint32_t c = row(r0 + i, buf.data());
            std::copy(buf.data(), buf.data() + c, g->colidx.data() + g->rowptr[i]);
        }
    }
    *out = g;
    *n_padded = n;
    *nnz = (int64_t)g->rowptr[nl];
    return MGGCN_OK;
}

// Rows [r0, r1) of the generated graph (r1 < 0: all of them).  The edge samples are counter-based, so a rank that
// wants only its row block walks the same sample stream and keeps the pairs that touch its rows: the rows it builds
// are entry for entry the rows of the whole graph.
static int gen_rows(int64_t n_raw, int64_t target_directed_edges, int degree_model, double gamma, int64_t max_deg,
                    uint64_t seed, int permute, int64_t r0, int64_t r1, mggcn_gen_t **out, int64_t *n_padded, int64_t *nnz) {
    if (n_raw <= 0 || target_directed_edges < 0 || !out || !n_padded || !nnz)
        return fail(MGGCN_E_INVALID, "bad argument");
    const int64_t n = (n_raw + 7) / 8 * 8;
    if (n > 0xFFFFFFFFll) return fail(MGGCN_E_UNSUPPORTED, "more than 2^32 vertices");
    if (r1 < 0) { r0 = 0; r1 = n; }
    if (r0 < 0 || r0 > r1 || r1 > n) return fail(MGGCN_E_INVALID, "row range [%lld, %lld) outside [0, %lld)", (long long)r0, (long long)r1, (long long)n);
    const int64_t nl = r1 - r0;
    const uint32_t lo32 = (uint32_t)r0;
    auto mine = [&](uint32_t x) { return (uint64_t)x - lo32 < (uint64_t)nl; };
    const int64_t E = target_directed_edges / 2;  // undirected samples; each yields two entries

    if (degree_model == 2) return gen_rows_affine(n_raw, n, target_directed_edges, seed, permute, r0, r1, out, n_padded, nnz);

    alias_table tab;
    if (degree_model == 1) {
        // Chung-Lu weights w_i = (i + i0)^(-a), a = 1/(gamma-1); i0 set so that the heaviest
        // vertex has expected degree max_deg.
        const double a = 1.0 / std::max(1e-3, gamma - 1.0);
        auto top_degree = [&](double i0) {
            // fixed 64Ki blocks summed in block order: the same value for any thread count.  Past 4 Mi vertices
            // (papers100M-shape: 60 bisection steps x 111 M pow() calls would take minutes) only the first 4 Mi
            // terms are summed and the smooth tail is integrated (midpoint rule, relative error < 1e-9 there);
            // graphs below that size are generated exactly as before.
            const int64_t n_sum = std::min<int64_t>(n_raw, 4ll << 20);
            double tail = 0;
            if (n_sum < n_raw) {
                const double x0 = (double)n_sum - 0.5 + i0, x1 = (double)n_raw - 0.5 + i0;
                tail = std::fabs(a - 1.0) < 1e-12 ? std::log(x1 / x0) : (std::pow(x1, 1.0 - a) - std::pow(x0, 1.0 - a)) / (1.0 - a);
            }
            const int64_t nb = (n_sum + 65535) / 65536;
            std::vector<double> part((size_t)nb);
#pragma omp parallel for schedule(static)
            for (int64_t b = 0; b < nb; ++b) {
                double s = 0;
                const int64_t hi_i = std::min<int64_t>(n_sum, (b + 1) * 65536);
                for (int64_t i = b * 65536; i < hi_i; ++i) s += std::pow((double)i + i0, -a);
                part[b] = s;
            }
            double W = tail;
            for (double s : part) W += s;
            return 2.0 * (double)E * std::pow(i0, -a) / W;
        };
        double lo = 1e-3, hi = (double)n_raw;
        if (max_deg <= 0) max_deg = n_raw / 10;
        for (int it = 0; it < 60; ++it) {
            double mid = std::sqrt(lo * hi);
            (top_degree(mid) > (double)max_deg ? lo : hi) = mid;
        }
        const double i0 = std::sqrt(lo * hi);
        std::vector<double> w((size_t)n_raw);
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n_raw; ++i) w[i] = std::pow((double)i + i0, -a);
        tab.build(w);
    } else if (degree_model != 0) {
        return fail(MGGCN_E_INVALID, "unknown degree model %d", degree_model);
    }
    const bool permuted = permute != 0;
    const std::vector<uint32_t> perm = permuted ? make_perm(n, seed ? seed : 1) : std::vector<uint32_t>();

    auto endpoints = [&](int64_t k, uint32_t &u, uint32_t &v) {
        uint64_t hu = hash2(seed, 2 * (uint64_t)k), hv = hash2(seed, 2 * (uint64_t)k + 1);
        if (degree_model == 1) {
            u = tab.sample(hu);
            v = tab.sample(hv);
        } else {
            u = (uint32_t)(((hu >> 32) * (uint64_t)n_raw) >> 32);
            v = (uint32_t)(((hv >> 32) * (uint64_t)n_raw) >> 32);
        }
        if (permuted) {
            u = perm[u];
            v = perm[v];
        }
    };

    const bool trace = getenv("MGGCN_GEN_TRACE") != nullptr;
    double t_prev = omp_get_wtime();
    auto lap = [&](const char *what) {
        if (!trace) return;
        const double t = omp_get_wtime();
        fprintf(stderr, "[gen] %-28s %.3f s\n", what, t - t_prev);
        t_prev = t;
    };
    lap("tables");
    // pass 1: one walk over the sample stream; the (local row, column) pairs that touch these rows go into
    // per-thread buffers and are counted.  Self pairs are dropped, the self-loop is added later.
    std::vector<std::atomic<uint32_t>> deg((size_t)nl);
    for (auto &d : deg) d.store(0, std::memory_order_relaxed);
    const int T = std::max(1, omp_get_max_threads());
    std::vector<std::vector<uint64_t>> kept((size_t)T);
#pragma omp parallel num_threads(T)
    {
        std::vector<uint64_t> &mine_buf = kept[(size_t)omp_get_thread_num()];
        mine_buf.reserve((size_t)((double)E * 2.2 * (double)nl / (double)n / T) + 1024);
#pragma omp for schedule(static)
        for (int64_t k = 0; k < E; ++k) {
            uint32_t u, v;
            endpoints(k, u, v);
            if (u == v) continue;
            if (mine(u)) {
                deg[u - lo32].fetch_add(1, std::memory_order_relaxed);
                mine_buf.push_back((uint64_t)(u - lo32) << 32 | v);
            }
            if (mine(v)) {
                deg[v - lo32].fetch_add(1, std::memory_order_relaxed);
                mine_buf.push_back((uint64_t)(v - lo32) << 32 | u);
            }
        }
    }
    lap("sample + keep");
    std::vector<uint64_t> start((size_t)nl + 1, 0);
    for (int64_t i = 0; i < nl; ++i) start[i + 1] = start[i] + deg[i].load(std::memory_order_relaxed);
    std::vector<uint32_t> raw(start[nl]);
    std::vector<std::atomic<uint64_t>> cursor((size_t)nl);
    for (int64_t i = 0; i < nl; ++i) cursor[i].store(start[i], std::memory_order_relaxed);
    // pass 2: scatter from the buffers (order inside a row is arbitrary here, fixed by the sort below)
    // (a loop over the T buffers, not "buffer of my thread id": this region may be given fewer threads than pass 1 had)
#pragma omp parallel for schedule(static, 1) num_threads(T)
    for (int t = 0; t < T; ++t) {
        std::vector<uint64_t> &mine_buf = kept[(size_t)t];
        for (uint64_t pr : mine_buf) raw[cursor[pr >> 32].fetch_add(1, std::memory_order_relaxed)] = (uint32_t)pr;
        std::vector<uint64_t>().swap(mine_buf);
    }
    lap("scatter");
    // pass 3: sort + dedup each row, insert the self-loop
    std::vector<uint32_t> final_deg((size_t)nl);
#pragma omp parallel for schedule(dynamic, 256)
    for (int64_t i = 0; i < nl; ++i) {
        uint32_t *b = raw.data() + start[i], *e = raw.data() + start[i + 1];
        std::sort(b, e);
        uint32_t *u = std::unique(b, e);
        final_deg[i] = (uint32_t)(u - b) + 1;  // + self-loop (self pairs were never stored)
    }
    auto *g = new (std::nothrow) mggcn_gen;
    if (!g) return fail(MGGCN_E_NOMEM, "out of memory");
    g->n = nl;
    g->rowptr.assign((size_t)nl + 1, 0);
    for (int64_t i = 0; i < nl; ++i) g->rowptr[i + 1] = g->rowptr[i] + final_deg[i];
    g->colidx.resize(g->rowptr[nl]);
#pragma omp parallel for schedule(dynamic, 256)
    for (int64_t i = 0; i < nl; ++i) {
        const uint32_t *b = raw.data() + start[i];
        const uint32_t cnt = final_deg[i] - 1;
        const uint32_t self = (uint32_t)(i + r0);
        uint32_t *o = g->colidx.data() + g->rowptr[i];
        uint32_t k = 0;
        while (k < cnt && b[k] < self) { *o++ = b[k]; ++k; }
        *o++ = self;
        while (k < cnt) { *o++ = b[k]; ++k; }
    }
    lap("sort + dedup + self-loops");
    *out = g;
    *n_padded = n;
    *nnz = (int64_t)g->rowptr[nl];
    return MGGCN_OK;
}

extern "C" {

int mggcn_gen_graph_begin(int64_t n_raw, int64_t target_directed_edges, int degree_model,
                          double gamma, int64_t max_deg, uint64_t seed, int permute,
                          mggcn_gen_t **out, int64_t *n_padded, int64_t *nnz) {
    return gen_rows(n_raw, target_directed_edges, degree_model, gamma, max_deg, seed, permute, 0, -1, out, n_padded, nnz);
}

int mggcn_gen_rows_begin(int64_t n_raw, int64_t target_directed_edges, int degree_model, double gamma,
                         int64_t max_deg, uint64_t seed, int permute, int64_t row_begin, int64_t row_end,
                         mggcn_gen_t **out, int64_t *n_padded, int64_t *nnz_rows) {
    if (row_end < 0) return fail(MGGCN_E_INVALID, "negative row range");
    return gen_rows(n_raw, target_directed_edges, degree_model, gamma, max_deg, seed, permute, row_begin, row_end, out,
                    n_padded, nnz_rows);
}

int mggcn_gen_graph_finish(mggcn_gen_t *g, uint64_t *rowptr, uint32_t *colidx, float *vals) {
    if (!g) return fail(MGGCN_E_INVALID, "null generator");
    if (rowptr) std::copy(g->rowptr.begin(), g->rowptr.end(), rowptr);
    const int64_t nnz = (int64_t)g->rowptr[g->n];
    if (colidx) {
#pragma omp parallel for schedule(static)
        for (int64_t e = 0; e < nnz; ++e) colidx[e] = g->colidx[e];
    }
    if (vals) {
#pragma omp parallel for schedule(static)
        for (int64_t e = 0; e < nnz; ++e) vals[e] = 1.0f;
    }
    delete g;
    return MGGCN_OK;
}

int mggcn_gen_features(int64_t n_raw, int64_t n_padded, int64_t f_raw, int64_t f_padded,
                       uint64_t seed, int permute, float *out) {
    if (!out || n_raw > n_padded || f_raw > f_padded) return fail(MGGCN_E_INVALID, "bad argument");
    const std::vector<uint32_t> perm = make_perm(n_padded, permute ? (seed ? seed : 1) : 0);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n_padded; ++i) {
        float *row = out + (size_t)perm[i] * f_padded;
        for (int64_t j = 0; j < f_padded; ++j) {
            if (i < n_raw && j < f_raw)
                row[j] = (float)(2.0 * u01(hash2(seed ^ 0x46454154ull, (uint64_t)i * f_raw + j)) - 1.0);
            else
                row[j] = 0.f;
        }
    }
    return MGGCN_OK;
}

// Rows [r0, r1) of what mggcn_gen_features / mggcn_gen_labels produce (out holds r1 - r0 rows).
int mggcn_gen_features_rows(int64_t n_raw, int64_t n_padded, int64_t f_raw, int64_t f_padded, uint64_t seed,
                            int permute, int64_t row_begin, int64_t row_end, float *out) {
    if (!out || n_raw > n_padded || f_raw > f_padded || row_begin < 0 || row_begin > row_end || row_end > n_padded)
        return fail(MGGCN_E_INVALID, "bad argument");
    std::vector<uint32_t> inv;  // old id of a new (permuted) id
    if (permute) {
        const std::vector<uint32_t> perm = make_perm(n_padded, seed ? seed : 1);
        inv.resize((size_t)n_padded);
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n_padded; ++i) inv[perm[i]] = (uint32_t)i;
    }
#pragma omp parallel for schedule(static)
    for (int64_t r = row_begin; r < row_end; ++r) {
        const int64_t i = permute ? (int64_t)inv[r] : r;
        float *row = out + (size_t)(r - row_begin) * f_padded;
        for (int64_t j = 0; j < f_padded; ++j) {
            if (i < n_raw && j < f_raw)
                row[j] = (float)(2.0 * u01(hash2(seed ^ 0x46454154ull, (uint64_t)i * f_raw + j)) - 1.0);
            else
                row[j] = 0.f;
        }
    }
    return MGGCN_OK;
}

// Feature rows of an arbitrary list of (permuted) vertex ids -- sampled checks need the rows of a vertex's
// neighbours, wherever they live.
int mggcn_gen_features_ids(int64_t n_raw, int64_t n_padded, int64_t f_raw, int64_t f_padded, uint64_t seed, int permute,
                           const uint32_t *ids, int64_t count, float *out) {
    if (!out || !ids || n_raw > n_padded || f_raw > f_padded || count < 0) return fail(MGGCN_E_INVALID, "bad argument");
    std::vector<uint32_t> inv;
    if (permute) {
        const std::vector<uint32_t> perm = make_perm(n_padded, seed ? seed : 1);
        inv.resize((size_t)n_padded);
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n_padded; ++i) inv[perm[i]] = (uint32_t)i;
    }
    for (int64_t k = 0; k < count; ++k)
        if ((int64_t)ids[k] >= n_padded) return fail(MGGCN_E_INVALID, "vertex id out of range");
#pragma omp parallel for schedule(static)
    for (int64_t k = 0; k < count; ++k) {
        const int64_t i = permute ? (int64_t)inv[ids[k]] : (int64_t)ids[k];
        float *row = out + (size_t)k * f_padded;
        for (int64_t j = 0; j < f_padded; ++j)
            row[j] = i < n_raw && j < f_raw ? (float)(2.0 * u01(hash2(seed ^ 0x46454154ull, (uint64_t)i * f_raw + j)) - 1.0) : 0.f;
    }
    return MGGCN_OK;
}

int mggcn_gen_labels_rows(int64_t n_raw, int64_t n_padded, int num_classes, uint64_t seed, int permute,
                          int64_t row_begin, int64_t row_end, int32_t *out) {
    if (!out || n_raw > n_padded || num_classes <= 0 || row_begin < 0 || row_begin > row_end || row_end > n_padded)
        return fail(MGGCN_E_INVALID, "bad argument");
    std::vector<uint32_t> inv;
    if (permute) {
        const std::vector<uint32_t> perm = make_perm(n_padded, seed ? seed : 1);
        inv.resize((size_t)n_padded);
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n_padded; ++i) inv[perm[i]] = (uint32_t)i;
    }
#pragma omp parallel for schedule(static)
    for (int64_t r = row_begin; r < row_end; ++r) {
        const int64_t i = permute ? (int64_t)inv[r] : r;
        out[r - row_begin] = i < n_raw ? (int32_t)(hash2(seed ^ 0x4C41424Cull, (uint64_t)i) % (uint64_t)num_classes) : 0;
    }
    return MGGCN_OK;
}

int mggcn_gen_labels(int64_t n_raw, int64_t n_padded, int num_classes, uint64_t seed, int permute,
                     int32_t *out) {
    if (!out || n_raw > n_padded || num_classes <= 0) return fail(MGGCN_E_INVALID, "bad argument");
    const std::vector<uint32_t> perm = make_perm(n_padded, permute ? (seed ? seed : 1) : 0);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n_padded; ++i)
        out[perm[i]] = i < n_raw ? (int32_t)(hash2(seed ^ 0x4C41424Cull, (uint64_t)i) % (uint64_t)num_classes) : 0;
    return MGGCN_OK;
}

int mggcn_init_weights(int64_t rows, int64_t cols, float gain, float *out) {
    if (!out || rows <= 0 || cols <= 0) return fail(MGGCN_E_INVALID, "bad argument");
    // same engine, seed and distribution types as dn_matrix::init (src/matrix.hpp:539-545) so the
    // libstdc++ stream is reproduced draw for draw
    // libstdc++'s uniform_real_distribution returns  canonical * (b - a) + a.  g++ contracts that into one fused
    // multiply-add whenever the target has FMA (the reference builds with -O3 -march=native, and oracle/_ref with
    // -march=x86-64-v3), so the reference binary's initial weights are the FMA results; written out here so the
    // stream is bit-identical to the reference build whatever flags THIS file is compiled with.
    std::default_random_engine gen(99);
    gain *= std::sqrt(3.0 / rows);
    const float a = -gain, b = gain;
    for (int64_t i = 0; i < rows * cols; ++i)
        out[i] = std::fmaf(std::generate_canonical<float, std::numeric_limits<float>::digits>(gen), b - a, a);
    return MGGCN_OK;
}

}  // extern "C"
