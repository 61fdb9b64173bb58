"""Pins oracle/ (the CPU restatement) on every known answer the reference's own tests hold for this
path: test/test_gcn.cpp:98-249 (loss, LeakyReLU, dense layer, and test_csr_g -- the only test through
cusparseSpMM), test/test_matrix.cpp:11-76,93-109 (toyA/toyB shapes, dense image, transpose) and the
weight-init stream of src/matrix.hpp:539-545.  Vectors live in tests/golden (tools/make_golden.py)."""
import json
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(__file__), "golden")
KA = json.load(open(os.path.join(GOLD, "known_answers.json")))
# the reference prints 7 significant digits; its own ASSERT_CLOSE allows 6.9e-5 relative
RTOL = 2e-6


def close(a, b, rtol=RTOL, atol=2e-7):
    np.testing.assert_allclose(np.asarray(a, np.float64), np.asarray(b, np.float64), rtol=rtol, atol=atol)


def test_cross_entropy(orc):
    k = KA["cross_entropy"]
    ls, cs, G = orc.softmax_xent(np.array(k["logits"], np.float32).reshape(3, 3), k["labels"])
    close(ls / 3, k["loss"])
    close(G.ravel(), k["G"])


def test_leaky_relu(orc):
    k = KA["leaky_relu"]
    logits = np.array(k["logits"], np.float32).reshape(3, 3)
    H = orc.leaky_relu_forward(logits)
    ls, cs, G = orc.softmax_xent(H, k["labels"])
    G = orc.leaky_relu_backward(logits, G)
    close(ls / 3, k["loss"])
    close(G.ravel(), k["G"])


@pytest.mark.parametrize("sparse", [False, True])
def test_layer_known_answer(orc, sparse):
    """test_g (dense A) and test_csr_g (CSR A: SpMM with beta = 1 onto a bias-prefilled C, backward
    through A.transpose())."""
    k = KA["layer"]
    X = np.array(k["X"], np.float32).reshape(2, 3)
    W = np.array(k["W"], np.float32).reshape(3, 2)
    b = np.array(k["b"], np.float32)
    Y = np.array(k["Y"], np.int32)
    XW = orc.gemm(X, W)
    AXW = orc.broadcast_rows(b, 2)
    if sparse:
        c = k["A_csr"]
        rp, ci, va = np.array(c["indptr"], np.uint64), np.array(c["indices"], np.uint32), np.array(c["data"], np.float32)
        AXW = orc.spmm(rp, ci, va, XW, C_in=AXW, alpha=1.0, beta=1.0)
    else:
        A = np.array(k["A_dense"], np.float32).reshape(2, 2)
        AXW = orc.gemm(A, XW, beta=1.0, C_in=AXW)
    H = orc.leaky_relu_forward(AXW)
    ls, cs, G = orc.softmax_xent(H, Y)
    G = orc.leaky_relu_backward(AXW, G)
    G_b = orc.gemm(np.ones((1, 2), np.float32), G)
    if sparse:
        trp, tci, tva = orc.transpose(rp, ci, va, 2)
        G_XW = orc.spmm(trp, tci, tva, G)
    else:
        G_XW = orc.gemm(A, G, ta=True)
    G_W = orc.gemm(X, G_XW, ta=True)
    G_out = orc.gemm(G_XW, W, tb=True)
    close(ls / 2, k["loss"])
    close(G.ravel(), k["G"])
    close(G_b.ravel(), k["G_b"])
    close(G_W.ravel(), k["G_W"])
    close(G_out.ravel(), k["G_out"], rtol=1e-5)  # the reference prints 0.4873930 vs 0.4873929: 7th digit


def _read_toy(name):
    raw = open(os.path.join(GOLD, name, "graph.bin"), "rb").read()
    assert raw[:11] == b"PIGO-CSR-v2" and raw[11] == 4 and raw[12] == 4
    n, m, nr, nc = np.frombuffer(raw, np.uint32, 4, 13)
    off = 13 + 16
    rp = np.frombuffer(raw, np.uint32, n + 1, off).astype(np.uint64)
    ci = np.frombuffer(raw, np.uint32, m, off + 4 * (n + 1)).copy()
    va = np.frombuffer(raw, np.float32, m, off + 4 * (n + 1) + 4 * m).copy()
    return rp, ci, va, int(nc)


def test_toy_dense_image_and_transpose(orc):
    rp, ci, va, m = _read_toy("toyA")
    assert (len(rp) - 1, m, len(ci)) == (KA["toyA"]["n"], KA["toyA"]["m"], KA["toyA"]["nnz"])
    D = orc.csr_to_dense(rp, ci, va, m)
    assert D.ravel().tolist() == KA["toyA"]["dense"]          # test_matrix.cpp:64-76
    trp, tci, tva = orc.transpose(rp, ci, va, m)
    assert np.array_equal(orc.csr_to_dense(trp, tci, tva, 4), D.T)  # test_matrix.cpp:93-109
    rpB, ciB, vaB, mB = _read_toy("toyB")
    assert (len(rpB) - 1, mB, len(ciB)) == (4, 4, KA["toyB"]["nnz"])
    # transpose output rows are sorted (serial form, src/matrix.hpp:426-453)
    t = orc.transpose(rpB, ciB, vaB, 4)
    for r in range(4):
        seg = t[1][int(t[0][r]):int(t[0][r + 1])]
        assert np.all(np.diff(seg.astype(np.int64)) > 0)


def test_init_stream(orc):
    assert np.array_equal(orc.init_weights(4, 3).ravel(), np.array(KA["init"]["W_4x3_default_gain"], np.float32))
    assert np.array_equal(orc.init_weights(1, 5, gain=np.sqrt(np.float32(1.0) / 3)).ravel(),
                          np.array(KA["init"]["b_1x5_gain_sqrt_third"], np.float32))


def test_normalize_unit_weights_is_inverse_indegree(orc):
    """Appendix B.1: with unit weights the column sum is an exact integer, so the result is 1/indeg(col)."""
    rp, ci, va, m = _read_toy("toyB")
    out = orc.normalize(rp, ci, va, m, axis=True)
    indeg = np.bincount(ci, minlength=m).astype(np.float32)
    assert np.array_equal(out, (np.float32(1) / indeg[ci]).astype(np.float32))


def test_staged_spmm_equals_blockwise_sum(orc):
    """orc_spmm with cuts restates src/cuda_utils.hpp:57-92: stage partials added in block order."""
    rng = np.random.default_rng(0)
    n, F, P = 64, 12, 4
    dense = (rng.random((n, n)) < 0.2) * rng.standard_normal((n, n))
    rp = np.zeros(n + 1, np.uint64)
    ci, va = [], []
    for r in range(n):
        nz = np.nonzero(dense[r])[0]
        ci += nz.tolist(); va += dense[r, nz].tolist()
        rp[r + 1] = len(ci)
    ci, va = np.array(ci, np.uint32), np.array(va, np.float32)
    B = rng.standard_normal((n, F)).astype(np.float32)
    cuts = np.array([i * n // P for i in range(P + 1)], np.uint32)
    got = orc.spmm(rp, ci, va, B, cuts=cuts)
    want = np.zeros((n, F), np.float32)
    for i in range(P):
        brp, bci, bva = orc.block(rp, ci, va, 0, n, int(cuts[i]), int(cuts[i + 1]))
        part = orc.spmm(brp, bci, bva, B[cuts[i]:cuts[i + 1]])
        want = part if i == 0 else (want + part).astype(np.float32)
    assert np.array_equal(got, want)
    ref64 = orc.spmm(rp, ci, va, B, acc64=True)
    np.testing.assert_allclose(got, ref64, rtol=1e-5, atol=1e-5)
