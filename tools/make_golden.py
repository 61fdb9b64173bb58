"""Regenerates tests/golden/ from the reference checkout (run in the build container only).

  * toyA / toyB dataset directories: byte copies of the reference's own fixtures
    (/root/reference/test/data/toy{A,B}/*.bin) -- data files, the only inputs its tests read;
  * known_answers.json: the known-answer vectors typed into the reference's tests
    (test/test_gcn.cpp:98-249, test/test_matrix.cpp:64-76) and the weight-init draws measured with
    the reference's own dn_matrix::init recipe (src/matrix.hpp:539-545) compiled here.
"""
import json
import os
import shutil
import subprocess
import tempfile

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def init_draws(rows, cols, gain_expr):
    # the reference's init body, verbatim semantics: default_random_engine(99), gain *= sqrt(3/rows)
    src = r"""
#include <random>
#include <cmath>
#include <cstdio>
int main() {
    std::default_random_engine gen(99);
    float gain = %s;
    const std::size_t N = %d, M = %d;
    gain *= std::sqrt(3.0 / N);
    std::uniform_real_distribution<float> uni(-gain, gain);
    for (std::size_t i = 0; i < N * M; i++) printf("%%.9g\n", uni(gen));
}
""" % (gain_expr, rows, cols)
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "a.cpp"), "w").write(src)
        # same flags as the reference build in oracle/Makefile (the reference's CMake uses -O3 -march=native): on an
        # FMA target g++ contracts the distribution's  canonical * (b - a) + a  into one fused multiply-add, and those
        # are the values the reference binary starts from (checked against its W / b dumps, tests/test_ref_layers.py)
        subprocess.check_call(["g++", "-std=c++20", "-O3", "-march=x86-64-v3", "-o", os.path.join(d, "a"), os.path.join(d, "a.cpp")])
        return [float(x) for x in subprocess.check_output([os.path.join(d, "a")]).split()]


def main():
    os.makedirs(OUT, exist_ok=True)
    for toy in ("toyA", "toyB"):
        dst = os.path.join(OUT, toy)
        os.makedirs(dst, exist_ok=True)
        for f in ("graph.bin", "features.bin", "labels.bin", "sets.bin"):
            shutil.copyfile(os.path.join(REF, "test", "data", toy, f), os.path.join(dst, f))
    ka = {
        "source": "GT-TDAlab/MG-GCN test/test_gcn.cpp:98-249, test/test_matrix.cpp:11-76",
        "cross_entropy": {  # test_gcn.cpp:98-116
            "logits": [2, 1, 2, 4, 2, 1, 1, -1, 0], "labels": [0, 0, 1], "loss": 1.146482,
            "G": [-0.1925604, 0.0517875, 0.1407729, -0.0520684, 0.0380651, 0.0140034, 0.2217470, -0.3033231, 0.0815762]},
        "leaky_relu": {  # test_gcn.cpp:119-139
            "logits": [2, 1, 2, 4, 2, 1, 1, -1, 0], "labels": [0, 0, 1], "loss": 0.8637248,
            "G": [-0.1925604, 0.0517875, 0.1407729, -0.0520684, 0.0380651, 0.0140034, 0.1924448, -0.0026324, 0.0007080]},
        "layer": {  # test_g :141-193 (dense A) and test_csr_g :195-249 (CSR A through cusparseSpMM)
            "A_dense": [1, 0, 0.5, 0.5],
            "A_csr": {"indptr": [0, 1, 3], "indices": [0, 0, 1], "data": [1, 0.5, 0.5], "m": 2},
            "X": [4, 2, 1, 1, -1, 0], "W": [1, 2, -1, 0, 0.5, 1.5], "b": [1, 0.5], "Y": [0, 1],
            "loss": 3.2750449,
            "G": [-0.4992494, 0.4992494, 0.0237129, -0.0237129],
            "G_b": [-0.4755365, 0.4755365],
            "G_W": [-1.9377153, 1.9377153, -0.9866424, 0.9866424, -0.4873929, 0.4873929],
            "G_out": [0.4873929, 0.4873929, 0.4873930, -0.0118565, -0.0118565, -0.0118565]},
        "toyA": {"n": 4, "m": 4, "nnz": 8,  # test_matrix.cpp:11-27,64-76
                 "dense": [0, 1, 0, 1, 1, 0, 1, 0, 0, 1, 0, 1, 1, 0, 1, 0], "features_shape": [4, 2]},
        "toyB": {"n": 4, "m": 4, "nnz": 12},
        "reddit_prep_shape": {"n": 232968, "nnz": 114848860, "features": [232968, 608]},  # :45-62
        "init": {
            "W_4x3_default_gain": init_draws(4, 3, "std::sqrt(2 / (1 + 0.01 * 0.01))"),
            "b_1x5_gain_sqrt_third": init_draws(1, 5, "std::sqrt((float)1.0 / 3)"),
        },
    }
    json.dump(ka, open(os.path.join(OUT, "known_answers.json"), "w"), indent=1)
    print("wrote", OUT)


if __name__ == "__main__":
    main()
