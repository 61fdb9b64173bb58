// Tests of the C++ surface.  `test_surface host <fixture dir>` needs no GPU (files, normalise,
// transpose, dense image: the checks of the reference's test/test_matrix.cpp:10-121);
// `test_surface gpu` replays the reference's known-answer tests test/test_gcn.cpp:98-249 through
// the same calls (matmul, broadcast_rows, leaky_relu_*, softmax_cross_entropy_loss) with a real
// relative check (the reference's ASSERT_CLOSE is log2-based and vacuous for negative values);
// `test_surface dist` needs >= 2 GPUs: scatter/gather round trip (test/test_dist_matrix.cpp:12-40),
// staged matmul == single-GPU matmul, dist_gcn loss == gcn loss at equal class padding.
#include <cmath>
#include <iostream>
#include <random>

#include "mg_gcn/gcn.hpp"

using namespace mg_gcn;

static int failures = 0;
#define EXPECT(cond)                                                                     \
    do {                                                                                 \
        if (!(cond)) {                                                                   \
            std::cerr << __FILE__ << ':' << __LINE__ << ": EXPECT failed: " #cond "\n";  \
            ++failures;                                                                  \
        }                                                                                \
    } while (0)
static bool close(double got, double want, double tol = 1e-5) {
    return std::fabs(got - want) <= tol * std::max(1.0, std::fabs(want));
}
#define EXPECT_CLOSE(got, want)                                                                             \
    do {                                                                                                    \
        if (!close((got), (want))) {                                                                        \
            std::cerr << __FILE__ << ':' << __LINE__ << ": got " << (got) << ", want " << (want) << "\n";  \
            ++failures;                                                                                     \
        }                                                                                                   \
    } while (0)

using csr_f = csr_matrix<unsigned, unsigned, float>;

static void host_tests(const std::string &dir) {
    // test/test_matrix.cpp:11-43: shapes of the fixtures
    csr_f A(dir + "/toyA/graph.bin"), B(dir + "/toyB/graph.bin");
    EXPECT(A.n() == 4 && A.m() == 4 && A.nnz() == 8);
    EXPECT(B.n() == 4 && B.m() == 4 && B.nnz() == 12);
    dn_matrix<float> X(dir + "/toyA/features.bin");
    EXPECT(X.n() == 4 && X.m() == 2);
    const float xe[8] = {0, 1, 1, 0, 0, 1, 1, 0};
    for (int i = 0; i < 8; ++i) EXPECT(X[i] == xe[i]);
    dn_matrix<std::int32_t> Y(dir + "/toyA/labels.bin");
    EXPECT(Y.n() == 4 && Y.m() == 1 && Y[1] == 1 && Y[2] == 0);
    // :64-76 dense image of toyA
    auto D = A.as_dn();
    const float de[16] = {0, 1, 0, 1, 1, 0, 1, 0, 0, 1, 0, 1, 1, 0, 1, 0};
    for (int i = 0; i < 16; ++i) EXPECT(D[i] == de[i]);
    // :93-109 transpose of the CSR == transpose of the dense image
    auto Bt = B.transpose().as_dn();
    auto Bd = B.as_dn();
    for (std::size_t i = 0; i < 4; ++i)
        for (std::size_t j = 0; j < 4; ++j) EXPECT((Bt[{i, j}] == Bd[{j, i}]));
    // normalize(true): every entry divided by its column sum (src/matrix.hpp:350-362)
    csr_f C({0, 1, 3}, {0, 0, 1}, {1.f, 0.5f, 0.5f}, 2);
    C.normalize(true);
    EXPECT_CLOSE(C[0].second, 1.0 / 1.5);
    EXPECT_CLOSE(C[1].second, 0.5 / 1.5);
    EXPECT_CLOSE(C[2].second, 1.0);
    csr_f R({0, 1, 3}, {0, 0, 1}, {1.f, 0.5f, 1.5f}, 2);
    R.normalize(false);
    EXPECT_CLOSE(R[1].second, 0.25);
    EXPECT_CLOSE(R[2].second, 0.75);
    // bad extension is a matrix_error (src/matrix.hpp:277-284)
    bool threw = false;
    try {
        csr_f bad(dir + "/toyA/graph.txt");
    } catch (const matrix_error &) {
        threw = true;
    }
    EXPECT(threw);
    // weight init = the reference's stream (first draws for in = 4)
    dn_matrix<float> W(4, 1);
    W.init();
    EXPECT_CLOSE(W[0], -1.22278583);
    EXPECT_CLOSE(W[1], -1.17020714);
    EXPECT_CLOSE(W[2], 0.746877909);
    EXPECT_CLOSE(W[3], -0.229351223);
}

static void loss_tests(const context &ctx) {
    dn_matrix<std::int32_t> Y(3, 1);
    Y.init({0, 0, 1});
    dn_matrix<float> logits(3, 3);
    logits.init({2, 1, 2, 4, 2, 1, 1, -1, 0});
    {  // test/test_gcn.cpp:98-116
        softmax_cross_entropy_loss<float, std::int32_t> loss_layer("0_");
        auto [loss, acc] = loss_layer(ctx, logits, Y);
        auto G = loss_layer.backward();
        ctx.sync();
        EXPECT_CLOSE(loss, 1.146482);
        EXPECT_CLOSE(acc, 2.0 / 3.0);  // rows 0 and 1 are classified correctly
        const double e[9] = {-0.1925604, 0.0517875, 0.1407729, -0.0520684, 0.0380651, 0.0140034, 0.2217470, -0.3033231, 0.0815762};
        for (int i = 0; i < 9; ++i) EXPECT_CLOSE(G[i], e[i]);
        EXPECT(logits[0] == 2.f);  // copy = true leaves the logits alone
    }
    {  // :119-139
        softmax_cross_entropy_loss<float, std::int32_t> loss_layer("0_");
        dn_matrix<float> H(3, 3);
        leaky_relu_forward(ctx, logits, H);
        auto [loss, acc] = loss_layer(ctx, H, Y);
        (void)acc;
        auto G = loss_layer.backward();
        leaky_relu_backward(ctx, logits, G, G);
        ctx.sync();
        EXPECT_CLOSE(loss, 0.8637248);
        const double e[9] = {-0.1925604, 0.0517875, 0.1407729, -0.0520684, 0.0380651, 0.0140034, 0.1924448, -0.0026324, 0.0007080};
        for (int i = 0; i < 9; ++i) EXPECT_CLOSE(G[i], e[i]);
    }
}

// test/test_gcn.cpp:141-249: one 2-node layer, dense (test_g) and through the CSR SpMM (test_csr_g)
static void layer_tests(const context &ctx, bool sparse) {
    csr_f A({0, 1, 3}, {0, 0, 1}, {1.f, 0.5f, 0.5f}, 2);
    dn_matrix<float> Ad(2, 2);
    Ad.init({1, 0, 0.5, 0.5});
    dn_matrix<float> X(2, 3), W(3, 2), b(1, 2);
    X.init({4, 2, 1, 1, -1, 0});
    W.init({1, 2, -1, 0, 0.5, 1.5});
    b.init({1, 0.5});
    dn_matrix<std::int32_t> Y(2, 1);
    Y.init({0, 1});
    dn_matrix<float> XW(2, 2), AXW(2, 2), H(2, 2), ones(1, 2), G_b(1, 2), G_XW(2, 2), G_W(3, 2), G_out(2, 3);
    matmul(ctx, X, W, XW, 1.f, 0.f);
    broadcast_rows(ctx, b, AXW);
    if (sparse) {
        auto ext = get_matmul_buffer(ctx, A, XW, AXW, 1.f, 0.f);
        matmul(ctx, A, XW, AXW, ext, 1.f, 1.f);
    } else {
        matmul(ctx, Ad, XW, AXW, 1.f, 1.f);
    }
    leaky_relu_forward(ctx, AXW, H);
    softmax_cross_entropy_loss<float, std::int32_t> loss_layer("0_");
    ones.init({1, 1});
    auto [loss, acc] = loss_layer(ctx, H, Y);
    (void)acc;
    auto G = loss_layer.backward();
    leaky_relu_backward(ctx, AXW, G, G);
    matmul(ctx, ones, G, G_b, 1.f, 0.f);
    if (sparse) {
        auto A_t = A.transpose();
        auto ext = get_matmul_buffer(ctx, A_t, G, G_XW, 1.f, 0.f);
        matmul(ctx, A_t, G, G_XW, ext, 1.f, 0.f);
    } else {
        matmul(ctx, Ad, G, G_XW, 1.f, 0.f, true);
    }
    matmul(ctx, X, G_XW, G_W, 1.f, 0.f, true);
    matmul(ctx, G_XW, W, G_out, 1.f, 0.f, false, true);
    ctx.sync();
    EXPECT_CLOSE(loss, 3.2750449);
    const double eg[4] = {-0.4992494, 0.4992494, 0.0237129, -0.0237129};
    for (int i = 0; i < 4; ++i) EXPECT_CLOSE(G[i], eg[i]);
    const double eb[2] = {-0.4755365, 0.4755365};
    for (int i = 0; i < 2; ++i) EXPECT_CLOSE(G_b[i], eb[i]);
    const double ew[6] = {-1.9377153, 1.9377153, -0.9866424, 0.9866424, -0.4873929, 0.4873929};
    for (int i = 0; i < 6; ++i) EXPECT_CLOSE(G_W[i], ew[i]);
    const double eo[6] = {0.4873929, 0.4873929, 0.4873930, -0.0118565, -0.0118565, -0.0118565};
    for (int i = 0; i < 6; ++i) EXPECT_CLOSE(G_out[i], eo[i]);
}

#include "test_surface_dist.inc"

int main(int argc, char **argv) {
    const std::string mode = argc > 1 ? argv[1] : "host";
    if (mode == "host") {
        host_tests(argc > 2 ? argv[2] : "tests/golden");
    } else if (mode == "gpu") {
        if (!have_device()) {
            std::cerr << "test_surface gpu: no CUDA device (there is no CPU fallback)\n";
            return 2;
        }
        context ctx(0);
        loss_tests(ctx);
        layer_tests(ctx, false);
        layer_tests(ctx, true);
        model_test(ctx);
        linear_test(ctx);
        layers_test(ctx, {64, 32});  // every layer transforms first
        layers_test(ctx, {96, 32});  // layer 0 aggregates first (in < out): quirk Q2 path
    } else if (mode == "dist") {
        int n = 0;
        if (cudaGetDeviceCount(&n) != cudaSuccess || n < 2) {
            std::cerr << "test_surface dist: needs 2 GPUs\n";
            return 77;
        }
        dist_tests();
    } else {
        std::cerr << "usage: test_surface host [fixture dir] | gpu | dist\n";
        return 2;
    }
    std::cout << "test_surface " << mode << ": " << (failures ? "FAILED" : "ok") << " (" << failures << " failures)\n";
    return failures ? 1 : 0;
}
