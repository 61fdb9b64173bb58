"""Parity of the dense kernels (GEMM + fused epilogues, column sums, element-wise, loss, Adam) with
the oracle through the C-ABI.  fp32; tolerance 1e-5 relative to the row scale (north_star)."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("ta,tb,M,N,K", [(0, 0, 1000, 128, 608), (0, 0, 517, 41, 128), (0, 0, 300, 256, 100),
                                         (0, 1, 900, 608, 128), (0, 1, 333, 100, 47), (1, 0, 608, 128, 5000),
                                         (1, 0, 128, 48, 20000), (1, 0, 8, 128, 3000), (1, 1, 64, 50, 70),
                                         (0, 0, 1, 1, 1), (0, 0, 130, 130, 17)])
def test_gemm_matches_fp64(mg, cuda, ta, tb, M, N, K):
    torch = cuda
    rng = np.random.default_rng(M + N + K)
    A = rng.standard_normal((K, M) if ta else (M, K)).astype(np.float32)
    B = rng.standard_normal((N, K) if tb else (K, N)).astype(np.float32)
    opA = A.T if ta else A
    opB = B.T if tb else B
    ref = opA.astype(np.float64) @ opB.astype(np.float64)
    scale = (np.abs(opA).astype(np.float64) @ np.abs(opB).astype(np.float64)).max(1, keepdims=True)
    C = mg.gemm(torch.from_numpy(A).cuda(), torch.from_numpy(B).cuda(), torch.empty((M, N), device="cuda"), ta=bool(ta), tb=bool(tb))
    err = np.abs(C.cpu().numpy() - ref) / np.maximum(scale, 1e-30)
    assert float(err.max()) <= 1e-5, float(err.max())


@pytest.mark.parametrize("split", [0, 1])
def test_gemm_weight_gradient_under_cancellation(mg, cuda, split):
    """G_W = H^T G at a reddit-sized K with inputs whose products cancel (the truncating hi/lo split drops
    one-sided terms, and a bar relative to sum |a||b| would hide what that does to a sum that nearly cancels).
    H is sign-mixed, G = (softmax - onehot) / n as the loss kernel produces it (every column sums to ~0 over the
    vertices).  Per entry against fp64, for both split modes: |err| <= 1e-5 * max(|entry|, rms of the result)."""
    torch = cuda
    rng = np.random.default_rng(11)
    K, M, N = 233_000, 128, 64
    H = rng.standard_normal((K, M)).astype(np.float32)
    Z = rng.standard_normal((K, N)) * 2
    P = np.exp(Z - Z.max(1, keepdims=True)); P /= P.sum(1, keepdims=True)
    P[np.arange(K), rng.integers(0, N, K)] -= 1.0
    G = (P / K).astype(np.float32)
    want = H.T.astype(np.float64) @ G.astype(np.float64)
    rms = float(np.sqrt((want ** 2).mean()))
    cancel = float((np.abs(H.T).astype(np.float64) @ np.abs(G).astype(np.float64)).mean() / np.abs(want).mean())
    old = mg.lib().mggcn_tune(b"gemm_split", split)
    try:
        got = mg.gemm(torch.from_numpy(H).cuda(), torch.from_numpy(G).cuda(), torch.empty((M, N), device="cuda"),
                      ta=True).cpu().numpy().astype(np.float64)
    finally:
        mg.lib().mggcn_tune(b"gemm_split", old)
    err = got - want
    ratio = float((np.abs(err) / np.maximum(np.abs(want), rms)).max())
    print(f"split={split} cancellation sum|ab|/|sum ab| = {cancel:.0f}x  max err/max(|entry|, rms) = {ratio:.2e}  "
          f"mean signed err / rms = {err.mean() / rms:+.2e}")
    assert ratio <= 1e-5, ratio


def test_gemm_epilogues_alpha_beta_strided(mg, cuda):
    torch = cuda
    rng = np.random.default_rng(1)
    M, N, K = 700, 41, 128
    A = rng.standard_normal((M, K)).astype(np.float32)
    B = rng.standard_normal((K, N)).astype(np.float32)
    C0 = rng.standard_normal((M, 44)).astype(np.float32)
    Z = rng.standard_normal((M, 44)).astype(np.float32)
    bias = rng.standard_normal(N).astype(np.float32)
    base = 0.5 * (A.astype(np.float64) @ B) + 2.0 * C0[:, :N]
    At, Bt, Zt, bt = (torch.from_numpy(x).cuda() for x in (A, B, Z, bias))
    for epi, want in ((None, base),
                      (mg.make_epilogue(mg.EPI_LEAKY_RELU, 0.01, bias=bt), np.where(base + bias > 0, base + bias, 0.01 * (base + bias))),
                      (mg.make_epilogue(mg.EPI_LEAKY_RELU_BWD, 0.01, aux=Zt), np.where(Z[:, :N] > 0, base, 0.01 * base))):
        Ct = torch.from_numpy(C0).cuda()
        mg.gemm(At, Bt, Ct[:, :N], alpha=0.5, beta=2.0, epi=epi)
        got = Ct.cpu().numpy()
        np.testing.assert_allclose(got[:, :N], want, rtol=0, atol=2e-4)
        assert np.array_equal(got[:, N:], C0[:, N:])


@pytest.mark.parametrize("tb,M,N,K", [(1, 1000, 128, 128), (0, 777, 128, 608), (1, 515, 64, 41)])
def test_gemm_vector_epilogue_paths(mg, cuda, tb, M, N, K):
    """The float4 epilogue of the tensor-core kernel: bias + LeakyReLU, LeakyReLU gradient with aux, beta*C."""
    torch = cuda
    rng = np.random.default_rng(7)
    A = rng.standard_normal((M, K)).astype(np.float32)
    B = rng.standard_normal((N, K) if tb else (K, N)).astype(np.float32)
    C0 = rng.standard_normal((M, N)).astype(np.float32)
    Z = rng.standard_normal((M, N)).astype(np.float32)
    bias = rng.standard_normal(N).astype(np.float32)
    prod = A.astype(np.float64) @ (B.T if tb else B).astype(np.float64)
    tol = 1e-5 * float((np.abs(A).astype(np.float64) @ np.abs(B.T if tb else B)).max())
    At, Bt, Zt, bt = (torch.from_numpy(x).cuda() for x in (A, B, Z, bias))
    for epi, beta, want in ((mg.make_epilogue(mg.EPI_LEAKY_RELU, 0.01, bias=bt), 0.0, np.where(prod + bias > 0, prod + bias, 0.01 * (prod + bias))),
                            (mg.make_epilogue(mg.EPI_LEAKY_RELU_BWD, 0.01, aux=Zt), 0.0, np.where(Z > 0, prod, 0.01 * prod)),
                            (None, 1.0, prod + C0)):
        Ct = torch.from_numpy(C0.copy()).cuda()
        mg.gemm(At, Bt, Ct, tb=bool(tb), beta=beta, epi=epi)
        np.testing.assert_allclose(Ct.cpu().numpy(), want, rtol=0, atol=tol)


def test_colsum_and_elementwise(mg, orc, cuda):
    torch = cuda
    rng = np.random.default_rng(2)
    G = rng.standard_normal((5003, 47)).astype(np.float32)
    got = mg.colsum(torch.from_numpy(G).cuda()).cpu().numpy()
    np.testing.assert_allclose(got, G.astype(np.float64).sum(0), rtol=0, atol=5e-4)
    x = rng.standard_normal((333, 17)).astype(np.float32)
    g = rng.standard_normal((333, 17)).astype(np.float32)
    xt, gt = torch.from_numpy(x).cuda(), torch.from_numpy(g).cuda()
    out = torch.empty_like(xt)
    mg.leaky_relu_forward(xt, out)
    assert np.array_equal(out.cpu().numpy(), orc.leaky_relu_forward(x))
    mg.leaky_relu_backward(xt, gt, out)
    assert np.array_equal(out.cpu().numpy(), orc.leaky_relu_backward(x, g))
    row = rng.standard_normal(17).astype(np.float32)
    mg.broadcast_rows(torch.from_numpy(row).cuda(), out)
    assert np.array_equal(out.cpu().numpy(), orc.broadcast_rows(row, 333))


@pytest.mark.parametrize("rows,C", [(3, 3), (1000, 41), (777, 48), (64, 172), (10, 1), (50, 700)])
def test_softmax_xent_matches_oracle(mg, orc, cuda, rows, C):
    torch = cuda
    rng = np.random.default_rng(rows + C)
    Z = (rng.standard_normal((rows, C)) * 3).astype(np.float32)
    Y = rng.integers(0, C, rows).astype(np.int32)
    ls, cs, G = orc.softmax_xent(Z, Y, n_global=rows * 2)
    ld = (C + 3) // 4 * 4
    Zt = torch.zeros((rows, ld), device="cuda")
    Zt[:, :C] = torch.from_numpy(Z).cuda()
    got_ls, got_cs = mg.softmax_xent(Zt[:, :C], torch.from_numpy(Y).cuda(), n_global=rows * 2)
    # the reference build uses fast-math exp/log/divide (CMakeLists.txt:57): ~2 ulp from libm
    assert abs(got_ls - ls) <= 1e-5 * max(1.0, abs(ls))
    assert got_cs == cs
    np.testing.assert_allclose(Zt[:, :C].cpu().numpy(), G, rtol=1e-5, atol=1e-7 / rows)


def test_loss_known_answers(mg, cuda):
    torch = cuda
    ka = json.load(open(os.path.join(GOLD, "known_answers.json")))
    k = ka["cross_entropy"]
    Z = torch.tensor(k["logits"], dtype=torch.float32, device="cuda").reshape(3, 3)
    ls, _ = mg.softmax_xent(Z, torch.tensor(k["labels"], dtype=torch.int32, device="cuda"))
    np.testing.assert_allclose(ls / 3, k["loss"], rtol=1e-5)
    np.testing.assert_allclose(Z.cpu().numpy().ravel(), k["G"], rtol=1e-5, atol=1e-7)


def test_adam_matches_oracle(mg, orc, cuda):
    torch = cuda
    rng = np.random.default_rng(4)
    n, nw = 5000, 4000
    p = rng.standard_normal(n).astype(np.float32)
    m = np.zeros(n, np.float32); v = np.zeros(n, np.float32)
    pt, mt, vt = (torch.from_numpy(x.copy()).cuda() for x in (p, m, v))
    for step in range(1, 6):
        g = rng.standard_normal(n).astype(np.float32) * 0.1
        gt = torch.from_numpy(g.copy()).cuda()
        orc.adam(p[:nw], g[:nw], m[:nw], v[:nw], 1e-2, 0.9, 0.999, 5e-4, 1e-8, step)
        orc.adam(p[nw:], g[nw:], m[nw:], v[nw:], 1e-2, 0.9, 0.999, 0.0, 1e-8, step)
        mg.adam(pt, gt, mt, vt, 1e-2, 0.9, 0.999, 5e-4, 1e-8, step, n_decay=nw)
        np.testing.assert_allclose(pt.cpu().numpy(), p, rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(gt.cpu().numpy(), g, rtol=1e-6, atol=1e-7)
