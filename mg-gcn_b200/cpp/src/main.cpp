// mg_gcn -- command line of the reference (src/main.cpp:47-196), same options and output formats:
//
//   mg_gcn [-h] [-P gpus] [-R nonzero] [-E epochs] [-S x] [-N] train <dir> <L> <d1..dL>
//
//   -P n   number of GPUs (default 1)
//   -R r   non-zero enables the row partition; with -P > 1 and -R 0 nothing is trained, exactly
//          like the reference (its column-partition branch is commented out, src/main.cpp:171-189)
//   -E n   epochs (default 20)
//   -S x   takes a dummy argument; runs the exchange on the compute stream (no overlap)
//   -N     accepted, sets overlap back on (the reference's no_wait flag is never read)
// stderr: "<n> <nnz>", "num_labels = ..", "feature size = ..", then "<epoch> <loss> <acc> <seconds>" per epoch.
// CSV:    csvs/[permuted_]<name>_<sizes..>_<P>.csv relative to the cwd (written if the directory exists).
//
// Additions (long options, so the short ones keep their meaning):
//   --timers   run eagerly with the reference's named per-stage timers and dump them to the CSV
//              every epoch (the default replays one CUDA graph per epoch and writes no per-stage rows)
//   --exact    exact gradients instead of the reference's quirks Q1/Q2
//   --eager    separate train_forward / backward / adam_update calls, no graph
#include <getopt.h>

#include <chrono>
#include <fstream>
#include <iostream>

#include "mg_gcn/gcn.hpp"

using namespace mg_gcn;

namespace {

class arg_error : public std::runtime_error {
   public:
    template <typename T>
    explicit arg_error(T t) : std::runtime_error(t) {}
};

int usage_(const char *prog) {
    std::cout << "Usage: " << prog << " [-h] [-P num] [-R nonzero] [-E epochs] [-S x] [-N] [--timers] [--exact] [--eager] "
                 "train <dir> <L> <d1..dL>"
              << std::endl;
    return 0;
}
int help_() {
    std::cout << "\nMG-GCN (B200-native build): multi-GPU full-batch GCN training\n\n"
                 "Options:\n"
                 "    -P : number of GPUs\n"
                 "    -R : non-zero enables the row partition (required with -P > 1)\n"
                 "    -E : number of epochs (default 20)\n"
                 "    -S : (takes an argument) disable the overlap of exchange and SpMM\n"
                 "    --timers / --exact / --eager : see the header of src/main.cpp\n\n"
                 "Arguments:\n"
                 "    train <dir> <L> <d1..dL> : dataset directory, number of hidden layers, their widths\n";
    return EXIT_SUCCESS;
}

using x_t = unsigned;
using v_t = unsigned;
using r_t = float;

std::string csv_name(const std::filesystem::path &dir, const std::vector<std::size_t> &sizes, std::size_t P) {
    std::string name;
    bool permuted = false;
    for (const auto &part : (dir / "graph.bin").parent_path()) {
        if (part == "permuted") permuted = true;
        else if (!part.empty() && part != "/") name = (permuted ? std::string("permuted_") : std::string()) + part.string();
    }
    for (auto s : sizes) name += "_" + std::to_string(s);
    return "csvs/" + name + "_" + std::to_string(P) + ".csv";
}

template <typename step_t>
void epoch_loop(std::size_t epochs, step_t step) {
    for (std::size_t e = 0; e < epochs; ++e) {
        const auto start = std::chrono::system_clock::now();
        const std::pair<r_t, r_t> la = step(e);
        const double secs = std::chrono::duration<double>(std::chrono::system_clock::now() - start).count();
        std::cerr << e << ' ' << la.first << ' ' << la.second << ' ' << secs << std::endl;
    }
}

int main_(int argc, char **argv) {
    opterr = 0;
    std::size_t P = 1, row_partition = 0, num_epochs = 20;
    bool overlap = true, timers = false, exact = false, eager = false;
    const option longopts[] = {{"timers", no_argument, nullptr, 1000}, {"exact", no_argument, nullptr, 1001},
                               {"eager", no_argument, nullptr, 1002}, {nullptr, 0, nullptr, 0}};
    int c;
    while ((c = getopt_long(argc, argv, "+h?P:R:E:S:N", longopts, nullptr)) != -1) {
        switch (c) {
            case '?': return usage_(argv[0]);
            case 'h': usage_(argv[0]); return help_();
            case 'P': P = std::stoull(optarg); break;
            case 'R': row_partition = std::stoull(optarg); break;
            case 'E': num_epochs = std::stoull(optarg); break;
            case 'S': overlap = false; break;
            case 'N': overlap = true; break;
            case 1000: timers = true; break;
            case 1001: exact = true; break;
            case 1002: eager = true; break;
            default: throw arg_error("Unknown argument.");
        }
    }
    gcn_options opt;
    opt.exact_grads = exact;
    opt.timers = timers;
    opt.use_graph = !eager && !timers;

    while (optind < argc && argv[optind] != nullptr) {
        const std::string command = argv[optind++];
        if (command.compare(0, 5, "train") != 0) throw arg_error("Unknown command.");
        if (optind + 1 >= argc) throw arg_error("train needs <dir> <L> <d1..dL>");
        const std::filesystem::path dir = argv[optind++];

        csr_matrix<x_t, v_t, r_t> A(dir / "graph.bin");
        dn_matrix<r_t> X(dir / "features.bin");
        dn_matrix<std::int32_t> Y(dir / "labels.bin");
        dn_matrix<std::int32_t> S(dir / "sets.bin");  // loaded, never used: the reference trains on all rows

        std::cerr << A.n() << ' ' << A.nnz() << std::endl;
        const auto num_labels = 1 + *std::max_element(Y.begin(), Y.end());
        std::cerr << "num_labels = " << num_labels << std::endl;
        std::cerr << "feature size = " << X.m() << std::endl;

        const int num_sizes = std::stoi(argv[optind++]);
        if (optind + num_sizes > argc) throw arg_error("fewer layer widths than announced");
        std::vector<std::size_t> sizes{X.m()};
        for (int i = 0; i < num_sizes; ++i) sizes.push_back(std::stoull(argv[optind++]));
        sizes.push_back((std::size_t)num_labels);

        std::ofstream of(csv_name(dir, sizes, P));  // silently unwritable when csvs/ is missing, like the reference
        const r_t lr = 1e-2f, b1 = 0.9f, b2 = 0.999f, wd = 5e-4f, eps = 1e-8f;  // src/main.cpp:126,163

        if (P <= 1) {
            context ctx(0);
            gcn<x_t, v_t, r_t> G(A, sizes, false, opt);
            ctx.sync();
            epoch_loop(num_epochs, [&](std::size_t e) {
                std::pair<r_t, r_t> la;
                if (opt.use_graph) {
                    la = G.train_epoch(ctx, X, Y, lr, b1, b2, wd, eps);
                } else {
                    la = G.train_forward(ctx, X, Y);
                    G.backward(ctx);
                    G.adam_update(ctx, lr, b1, b2, wd, eps);
                }
                G.sync();
                if (timers) G.dump_timers(of, std::to_string(e) + "_0_");
                return la;
            });
        } else {
            sizes.back() = (sizes.back() + P - 1) / P * P;  // src/main.cpp:135
            dist_context ctx(P, overlap);
            std::vector<v_t> p(P + 1, 0);
            for (std::size_t i = 1; i < p.size(); ++i) p[i] = (v_t)((std::uint64_t)i * A.n() / P);
            A.normalize(true);
            auto A_T = A.transpose();
            if (row_partition) {
                dist_row_dn_matrix<std::int32_t> Yd(ctx, Y);
                dist_row_csr_matrix<x_t, v_t, r_t> Ad(ctx, A, p, p), A_Td(ctx, A_T, p, p);
                dist_gcn<true, x_t, v_t, r_t> G(ctx, Ad, A_Td, sizes, false, opt);
                dist_row_dn_matrix<r_t> Xd(ctx, X);
                ctx.sync();
                epoch_loop(num_epochs, [&](std::size_t e) {
                    std::pair<r_t, r_t> la;
                    if (opt.use_graph) {
                        la = G.train_epoch(ctx, Xd, Yd, lr, b1, b2, wd, eps);
                    } else {
                        la = G.train_forward(ctx, Xd, Yd);
                        G.backward(ctx);
                        G.adam_update(ctx, lr, b1, b2, wd, eps);
                    }
                    G.sync();
                    if (timers) G.dump_timers(of, std::to_string(e) + "_");
                    return la;
                });
            }
        }
    }
    return EXIT_SUCCESS;
}

}  // namespace

int main(int argc, char **argv) {
    // One thread feeds every GPU, and a GPU that has been launched spins in its stage barrier until its
    // peers are launched: nothing on the launch path may wait for a device, so kernels are loaded up
    // front instead of on first use (lazy loading allocates, and allocation waits for peer GPUs).
    setenv("CUDA_MODULE_LOADING", "EAGER", 0);
    try {
        return main_(argc, argv);
    } catch (const std::exception &e) {
        std::cerr << "Error: uncaught exception " << e.what() << std::endl;
        return EXIT_FAILURE;
    }
}
