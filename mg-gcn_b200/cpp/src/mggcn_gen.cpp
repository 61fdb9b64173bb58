// mggcn_gen -- writes one synthetic dataset directory (graph.bin, features.bin, labels.bin, sets.bin) in the
// reference's on-disk formats (test/data/prep.py:46-76,101-126), from the same generator entry points the library
// exports (csrc/host_gen.cpp, csrc/host_io.cpp), linked STATICALLY: no libmggcn.so, no CUDA runtime.
// It exists for bench.py's reference arm: that arm times the reference's own program and must not map this
// repository's library into its process, yet it needs the very same generated files as the repo arm.
//
//   mggcn_gen <n_raw> <edges> <f_raw> <classes> <degree_model> <gamma> <max_deg> <seed> <permute> <out_dir>
// prints one JSON line {"n": .., "nnz": .., "F": .., "classes": ..} on stdout; exit code 0 on success.
#include <sys/stat.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "mggcn.h"

#define TRY(call)                                                                    \
    do {                                                                             \
        const int rc__ = (call);                                                     \
        if (rc__ != 0) {                                                             \
            std::fprintf(stderr, "mggcn_gen: %s failed (%d): %s\n", #call, rc__, mggcn_last_error()); \
            return 1;                                                                \
        }                                                                            \
    } while (0)

int main(int argc, char **argv) {
    if (argc != 11) {
        std::fprintf(stderr, "usage: mggcn_gen n_raw edges f_raw classes degree_model gamma max_deg seed permute out_dir\n");
        return 2;
    }
    const std::int64_t n_raw = std::atoll(argv[1]), edges = std::atoll(argv[2]), f_raw = std::atoll(argv[3]);
    const int classes = std::atoi(argv[4]), model = std::atoi(argv[5]);
    const double gamma = std::atof(argv[6]);
    const std::int64_t max_deg = std::atoll(argv[7]);
    const std::uint64_t seed = std::strtoull(argv[8], nullptr, 10);
    const int permute = std::atoi(argv[9]);
    const std::string dir = argv[10];
    mkdir(dir.c_str(), 0777);  // the parent exists (bench.py creates it); an existing directory is fine

    mggcn_gen_t *g = nullptr;
    std::int64_t n = 0, nnz = 0;
    TRY(mggcn_gen_graph_begin(n_raw, edges, model, gamma, max_deg, seed, permute, &g, &n, &nnz));
    {
        std::vector<std::uint64_t> rowptr((std::size_t)n + 1);
        std::vector<std::uint32_t> colidx((std::size_t)nnz);
        std::vector<float> vals((std::size_t)nnz);
        TRY(mggcn_gen_graph_finish(g, rowptr.data(), colidx.data(), vals.data()));
        TRY(mggcn_csr_file_write((dir + "/graph.bin").c_str(), n, n, rowptr.data(), colidx.data(), vals.data(), 4));
    }
    const std::int64_t f8 = (f_raw + 7) / 8 * 8;
    {
        std::vector<float> X((std::size_t)n * (std::size_t)f8);
        TRY(mggcn_gen_features(n_raw, n, f_raw, f8, seed, permute, X.data()));
        TRY(mggcn_dense_file_write((dir + "/features.bin").c_str(), n, f8, X.data(), 4));
    }
    {
        std::vector<std::int32_t> Y((std::size_t)n);
        TRY(mggcn_gen_labels(n_raw, n, classes, seed, permute, Y.data()));
        TRY(mggcn_dense_file_write((dir + "/labels.bin").c_str(), n, 1, Y.data(), 4));  // labels are >= 0: same bits as u32
        std::vector<std::uint32_t> sets((std::size_t)n, 0u);
        TRY(mggcn_dense_file_write((dir + "/sets.bin").c_str(), n, 1, sets.data(), 4));
    }
    std::printf("{\"n\": %lld, \"nnz\": %lld, \"F\": %lld, \"classes\": %d}\n", (long long)n, (long long)nnz, (long long)f8, classes);
    return 0;
}
