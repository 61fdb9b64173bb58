// TEST INFRASTRUCTURE ONLY -- never linked into the product.
//
// Per-layer dumper over the REFERENCE's own classes ("link a small dumper against the reference
// headers").  Compiled by oracle/Makefile (target ref_dump) against /root/reference/src where it lies, next to
// mg_gcn_ref; it constructs gcn / dist_gcn<true> exactly like /root/reference/src/main.cpp:117-151 does, runs ONE
// forward + backward from the reference's initialisation and writes, in the reference's dense blob format
// (u32 N, u32 M, data):
//     H<l>.bin   output of layer l (operator() of the layer objects gcn::layers() hands out, src/gcn.hpp:437-452,
//                556-571) -- taken during a first forward pass, because backward overwrites the buffers in place
//     W<l>.bin b<l>.bin GW<l>.bin Gb<l>.bin   through the accessors src/gcn.hpp:503-517,622-635 after backward
//     loss.txt   the loss / accuracy train_forward returned
// north_star: "per-layer fp32 outputs and gradients must match the reference's cuSPARSE+NCCL build within 1e-5".
//
//   ref_dump <P> <dataset dir>/ <out dir> <L> <d1..dL>
#include <cstdint>
#include <filesystem>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "matrix.hpp"
#include "gcn.hpp"

using x_t = unsigned;
using v_t = unsigned;
using r_t = float;

static void write_blob(const std::filesystem::path &path, const std::vector<float> &host, std::size_t n, std::size_t m) {
    std::ofstream f(path, std::ios::binary);
    const std::uint32_t hdr[2] = {(std::uint32_t)n, (std::uint32_t)m};
    f.write(reinterpret_cast<const char *>(hdr), sizeof hdr);
    f.write(reinterpret_cast<const char *>(host.data()), (std::streamsize)(host.size() * sizeof(float)));
}

template <typename M>
static void append(std::vector<float> &host, const M &mat) {
    const std::size_t old = host.size();
    host.resize(old + mat.size());
    CHECK_CUDA( cudaMemcpy(host.data() + old, mat.buffer(), mat.size() * sizeof(float), cudaMemcpyDefault) );
}

int main(int argc, char **argv) {
    if (argc < 6) {
        std::cerr << "usage: ref_dump <P> <dataset dir>/ <out dir> <L> <d1..dL>" << std::endl;
        return 2;
    }
    const std::size_t P = std::stoull(argv[1]);
    const std::filesystem::path dir = argv[2], out = argv[3];
    std::filesystem::create_directories(out);

    csr_matrix<x_t, v_t, r_t> A(dir / "graph.bin");
    dn_matrix<r_t> X(dir / "features.bin");
    dn_matrix<std::int32_t> Y(dir / "labels.bin");
    const auto num_labels = 1 + *std::max_element(Y.begin(), Y.end());
    std::vector<std::size_t> sizes{X.m()};
    const int L = std::stoi(argv[4]);
    for (int i = 0; i < L; i++) sizes.push_back(std::stoull(argv[5 + i]));
    sizes.push_back(num_labels);
    const std::size_t nl = sizes.size() - 1;
    std::ofstream lossf(out / "loss.txt");

    if (P <= 1) {
        auto ctx = context(0);
        gcn<x_t, v_t, r_t> G(A, sizes);
        auto &layers = const_cast<std::vector<gcn_layer<x_t, v_t, r_t>> &>(G.layers());
        ctx.sync();
        dn_matrix<r_t> H = X;
        for (std::size_t l = 0; l < nl; l++) {
            H = layers[l](ctx, H);
            ctx.sync();
            std::vector<float> host;
            append(host, H);
            write_blob(out / ("H" + std::to_string(l) + ".bin"), host, H.n(), H.m());
        }
        auto [loss, acc] = G.train_forward(ctx, X, Y);
        G.backward(ctx);
        ctx.sync();
        lossf << loss << ' ' << acc << std::endl;
        for (std::size_t l = 0; l < nl; l++) {
            const std::string s = std::to_string(l);
            std::vector<float> w, b, gw, gb;
            append(w, layers[l].W());   write_blob(out / ("W" + s + ".bin"), w, layers[l].W().n(), layers[l].W().m());
            append(b, layers[l].b());   write_blob(out / ("b" + s + ".bin"), b, 1, layers[l].b().m());
            append(gw, layers[l].GW()); write_blob(out / ("GW" + s + ".bin"), gw, layers[l].GW().n(), layers[l].GW().m());
            append(gb, layers[l].Gb()); write_blob(out / ("Gb" + s + ".bin"), gb, 1, layers[l].Gb().m());
        }
    } else {
        sizes.back() = (sizes.back() + P - 1) / P * P;   // src/main.cpp:135
        auto ctx = dist_context(P, true);
        std::vector<v_t> p(P + 1);
        for (std::size_t i = 1; i < p.size(); i++) p[i] = i * A.n() / P;
        A.normalize(true);
        auto A_T = A.transpose();
        dist_row_dn_matrix<std::int32_t> Yd(ctx, Y);
        dist_row_csr_matrix Ad(ctx, A, p, p);
        dist_row_csr_matrix A_Td(ctx, A_T, p, p);
        dist_gcn<true, x_t, v_t, r_t> G(ctx, Ad, A_Td, sizes);
        dist_row_dn_matrix Xd(ctx, X);
        auto &layers = const_cast<std::vector<dist_gcn_layer<true, x_t, v_t, r_t>> &>(G.layers());
        ctx.sync();
        auto H = Xd;
        for (std::size_t l = 0; l < nl; l++) {
            H = layers[l](ctx, H);
            ctx.sync();
            std::vector<float> host;
            for (std::size_t i = 0; i < P; i++) append(host, H[i]);
            write_blob(out / ("H" + std::to_string(l) + ".bin"), host, A.n(), sizes[l + 1]);
        }
        auto [loss, acc] = G.train_forward(ctx, Xd, Yd);
        G.backward(ctx);
        ctx.sync();
        lossf << loss << ' ' << acc << std::endl;
        for (std::size_t l = 0; l < nl; l++) {
            const std::string s = std::to_string(l);
            std::vector<float> w, b, gw, gb;
            append(w, layers[l].W()[0]);   write_blob(out / ("W" + s + ".bin"), w, sizes[l], sizes[l + 1]);
            append(b, layers[l].b()[0]);   write_blob(out / ("b" + s + ".bin"), b, 1, sizes[l + 1]);
            append(gw, layers[l].GW()[0]); write_blob(out / ("GW" + s + ".bin"), gw, sizes[l], sizes[l + 1]);
            append(gb, layers[l].Gb()[0]); write_blob(out / ("Gb" + s + ".bin"), gb, 1, sizes[l + 1]);
        }
    }
    return 0;
}
