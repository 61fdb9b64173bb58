"""Parity of the sm_100a CSR SpMM (mggcn_spmm_csr_f32_*) with the oracle, through the C-ABI.

Tolerance (north_star): per-layer fp32 outputs within 1e-5 relative.  cuSPARSE's summation order is
unspecified, so the check is |got - ref64| <= 1e-5 * max(1, row-wise inf-norm scale), against the
fp64-accumulating oracle, and additionally <= a few ulp-sums against the fp32 oracle."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
TOL = 1e-5


def random_csr(rng, n, m, avg_deg, heavy=0, weights=True):
    deg = rng.poisson(avg_deg, n).astype(np.int64)
    deg = np.minimum(deg, m)
    if heavy:
        idx = rng.choice(n, size=min(heavy, n), replace=False)
        deg[idx] = np.minimum(m, rng.integers(3000, 9000, len(idx)))
    if n > 3:
        deg[1] = 0  # an empty row
    rp = np.zeros(n + 1, np.uint64)
    rp[1:] = np.cumsum(deg)
    ci = np.empty(int(rp[-1]), np.uint32)
    for r in range(n):
        if deg[r]:
            ci[int(rp[r]):int(rp[r + 1])] = np.sort(rng.choice(m, size=deg[r], replace=False))
    va = (rng.standard_normal(len(ci)).astype(np.float32)) if weights else np.ones(len(ci), np.float32)
    return rp, ci, va


def to_dev(torch, rp, ci, va, idx64=False):
    rpt = torch.from_numpy(rp.astype(np.int64) if idx64 else rp.astype(np.uint32).view(np.int32)).cuda()
    return rpt, torch.from_numpy(ci.view(np.int32)).cuda(), torch.from_numpy(va).cuda()


def check(got, ref64, scale_rows):
    err = np.abs(got.astype(np.float64) - ref64.astype(np.float64))
    scale = np.maximum(scale_rows, 1e-30)[:, None]
    assert float((err / scale).max()) <= TOL, float((err / scale).max())


@pytest.mark.parametrize("F,ld", [(8, 8), (41, 44), (47, 47), (64, 64), (100, 104), (112, 112), (128, 128),
                                  (172, 172), (256, 256), (300, 300), (1, 1), (33, 36)])
@pytest.mark.parametrize("use_plan", [False, True])
def test_spmm_matches_oracle(mg, orc, cuda, F, ld, use_plan):
    torch = cuda
    rng = np.random.default_rng(F * 2 + use_plan)
    n, m = 1500, 1300
    rp, ci, va = random_csr(rng, n, m, 40, heavy=3)
    B = rng.standard_normal((m, ld)).astype(np.float32)
    ref64 = orc.spmm(rp, ci, va, B, acc64=True, F=F)
    ref32 = orc.spmm(rp, ci, va, B, F=F)
    absrow = orc.spmm(rp, ci, np.abs(va), np.abs(B), acc64=True, F=F).max(1)
    rpt, cit, vat = to_dev(torch, rp, ci, va)
    Bt = torch.from_numpy(B).cuda()
    Ct = torch.full((n, ld), 7.0, dtype=torch.float32, device="cuda")  # beta = 0 must not read C
    plan = mg.SpmmPlan(rp.astype(np.uint32)) if use_plan else None
    mg.spmm(rpt, cit, vat, Bt, Ct, cols=m, F=F, plan=plan)
    got = Ct.cpu().numpy()
    check(got[:, :F], ref64, absrow)
    np.testing.assert_allclose(got[:, :F], ref32, rtol=0, atol=float(absrow.max()) * 2e-6)
    assert np.all(got[:, F:] == 7.0)  # padding columns of C untouched
    assert np.all(got[1, :F] == 0.0)  # empty row


@pytest.mark.parametrize("F", [41, 128])
def test_alpha_beta_epilogues_and_i64(mg, orc, cuda, F):
    torch = cuda
    rng = np.random.default_rng(11 + F)
    n = 700
    ld = (F + 3) // 4 * 4
    rp, ci, va = random_csr(rng, n, n, 25, heavy=2)
    B = rng.standard_normal((n, ld)).astype(np.float32)
    C0 = rng.standard_normal((n, ld)).astype(np.float32)
    Z = rng.standard_normal((n, ld)).astype(np.float32)
    bias = rng.standard_normal(F).astype(np.float32)
    absrow = orc.spmm(rp, ci, np.abs(va), np.abs(B), acc64=True, F=F).max(1) * 0.5 + 2 * np.abs(C0[:, :F]).max(1) + 2
    base = orc.spmm(rp, ci, va, B, C_in=C0[:, :F].copy(), alpha=0.5, beta=2.0, acc64=True, F=F)
    rpt, cit, vat = to_dev(torch, rp, ci, va, idx64=True)
    Bt, Zt, bt = torch.from_numpy(B).cuda(), torch.from_numpy(Z).cuda(), torch.from_numpy(bias).cuda()
    plan = mg.SpmmPlan(rp)  # uint64 row pointers
    cases = {
        "none": (None, base),
        "lrelu": (mg.make_epilogue(mg.EPI_LEAKY_RELU, 0.01), np.where(base > 0.01 * base, base, 0.01 * base)),
        "lrelu_bwd": (mg.make_epilogue(mg.EPI_LEAKY_RELU_BWD, 0.01, aux=Zt), np.where(Z[:, :F] > 0, base, 0.01 * base)),
        "bias_lrelu": (mg.make_epilogue(mg.EPI_LEAKY_RELU, 0.01, bias=bt),
                       np.where(base + bias > 0, base + bias, 0.01 * (base + bias))),
    }
    for name, (epi, want) in cases.items():
        for p in (None, plan):
            Ct = torch.from_numpy(C0).cuda()
            mg.spmm(rpt, cit, vat, Bt, Ct, cols=n, F=F, alpha=0.5, beta=2.0, epi=epi, plan=p)
            check(Ct.cpu().numpy()[:, :F], want, absrow)
    # dropout keep-mask epilogue (new functionality, off in the reference)
    mask = (rng.random((n, F)) < 0.7).astype(np.uint8)
    mt = torch.from_numpy(mask).cuda()
    Ct = torch.from_numpy(C0).cuda()
    mg.spmm(rpt, cit, vat, Bt, Ct, cols=n, F=F, alpha=0.5, beta=2.0,
            epi=mg.make_epilogue(mg.EPI_NONE, keep_mask=mt, keep_scale=1 / 0.7), plan=plan)
    check(Ct.cpu().numpy()[:, :F], np.where(mask, base / 0.7, 0.0), absrow / 0.7)


def test_degenerate_shapes(mg, cuda):
    torch = cuda
    B = torch.ones((5, 8), device="cuda")
    C = torch.full((3, 8), 3.0, device="cuda")
    rp = torch.zeros(4, dtype=torch.int32, device="cuda")  # all rows empty
    e = torch.zeros(1, dtype=torch.int32, device="cuda")
    v = torch.zeros(1, dtype=torch.float32, device="cuda")
    mg.spmm(rp, e, v, B, C, cols=5, F=8)
    assert float(C.abs().max()) == 0.0
    C.fill_(3.0)
    mg.spmm(rp, e, v, B, C, cols=5, F=8, beta=1.0)
    assert float((C - 3.0).abs().max()) == 0.0
    with pytest.raises(mg.MggcnError):  # ldc < F
        mg.spmm(rp, e, v, B, C, cols=5, F=16)
    rp0 = torch.zeros(1, dtype=torch.int32, device="cuda")  # zero rows: no-op
    mg.spmm(rp0, e, v, B, C[:0], cols=5, F=8)


def test_reference_known_answer_through_cuda(mg, cuda):
    """test/test_gcn.cpp:195-249 (test_csr_g) through the C-ABI: GEMM, SpMM with beta = 1 onto a
    bias-prefilled C, LeakyReLU, loss, LeakyReLU backward, column sums, transposed SpMM, G_W, G_out."""
    torch = cuda
    k = json.load(open(os.path.join(GOLD, "known_answers.json")))["layer"]
    f = lambda a, s: torch.tensor(a, dtype=torch.float32, device="cuda").reshape(s)
    X, W, b = f(k["X"], (2, 3)), f(k["W"], (3, 2)), f(k["b"], (2,))
    Y = torch.tensor(k["Y"], dtype=torch.int32, device="cuda")
    c = k["A_csr"]
    A = mg.Csr(np.array(c["indptr"]), np.array(c["indices"]), np.array(c["data"]), 2)
    dev = lambda M: (torch.from_numpy(M.rowptr.astype(np.uint32).view(np.int32)).cuda(),
                     torch.from_numpy(M.colidx.view(np.int32)).cuda(), torch.from_numpy(M.vals).cuda())
    XW = mg.gemm(X, W, torch.empty((2, 2), device="cuda"))
    AXW = torch.empty((2, 2), device="cuda")
    mg.broadcast_rows(b, AXW)
    mg.spmm(*dev(A), XW, AXW, beta=1.0)
    H = torch.empty_like(AXW)
    mg.leaky_relu_forward(AXW, H)
    loss_sum, _ = mg.softmax_xent(H, Y)
    G = H
    mg.leaky_relu_backward(AXW, G, G)
    G_b = mg.colsum(G)
    G_XW = mg.spmm(*dev(A.transpose()), G, torch.empty((2, 2), device="cuda"))
    G_W = mg.gemm(X, G_XW, torch.empty((3, 2), device="cuda"), ta=True)
    G_out = mg.gemm(G_XW, W, torch.empty((2, 3), device="cuda"), tb=True)
    close = lambda a, e: np.testing.assert_allclose(np.asarray(a, np.float64).ravel(), e, rtol=1e-5, atol=1e-7)
    close(loss_sum / 2, k["loss"])
    close(G.cpu().numpy(), k["G"])
    close(G_b.cpu().numpy(), k["G_b"])
    close(G_W.cpu().numpy(), k["G_W"])
    close(G_out.cpu().numpy(), k["G_out"])


def test_staged_blocks_match_oracle_stage_order(mg, orc, cuda):
    """The multi-GPU form on one device: C = sum_i A[.,i] * B_i with beta = 0 then 1
    (src/cuda_utils.hpp:82-83), block-local column ids, against orc_spmm with cuts."""
    torch = cuda
    rng = np.random.default_rng(5)
    n, F, P = 2000, 128, 4
    rp, ci, va = random_csr(rng, n, n, 60)
    A = mg.Csr(rp, ci, va, n)
    cuts = mg.row_cuts(n, P)
    B = rng.standard_normal((n, F)).astype(np.float32)
    want = orc.spmm(rp, ci, va, B, cuts=cuts)
    ref64 = orc.spmm(rp, ci, va, B, acc64=True)
    absrow = orc.spmm(rp, ci, np.abs(va), np.abs(B), acc64=True).max(1)
    Bt = torch.from_numpy(B).cuda()
    got = np.empty((n, F), np.float32)
    for i in range(P):
        blocks = A.block_rows(cuts, cuts, only_row=i)
        Ct = torch.empty((int(cuts[i + 1] - cuts[i]), F), device="cuda")
        for j, blk in enumerate(blocks):
            plan = mg.SpmmPlan(blk.rowptr.astype(np.uint32))
            mg.spmm(torch.from_numpy(blk.rowptr.astype(np.uint32).view(np.int32)).cuda(),
                    torch.from_numpy(blk.colidx.view(np.int32)).cuda(), torch.from_numpy(blk.vals).cuda(),
                    Bt[int(cuts[j]):int(cuts[j + 1])], Ct, cols=blk.m, F=F, beta=0.0 if j == 0 else 1.0, plan=plan)
        got[int(cuts[i]):int(cuts[i + 1])] = Ct.cpu().numpy()
    check(got, ref64, absrow)
    np.testing.assert_allclose(got, want, rtol=0, atol=float(absrow.max()) * 2e-6)


def test_full_size_properties_reddit_shape(mg, orc, cuda):
    """BASELINE.json config 2 at full size: size-independent properties + sampled rows vs the oracle."""
    torch = cuda
    A, _, _, _ = mg.generate("reddit", seed=1, features=False)
    A.normalize(by_column=True)
    AT = A.transpose()  # forward operand: rows sum to 1 (mean over in-neighbours)
    n, F = AT.n, 128
    rpt = torch.from_numpy(AT.rowptr.astype(np.uint32).view(np.int32)).cuda()
    cit = torch.from_numpy(AT.colidx.view(np.int32)).cuda()
    vat = torch.from_numpy(AT.vals).cuda()
    plan = mg.SpmmPlan(AT.rowptr.astype(np.uint32))
    ones = torch.ones((n, F), device="cuda")
    C1 = mg.spmm(rpt, cit, vat, ones, torch.empty((n, F), device="cuda"), plan=plan)
    # row-stochastic: A_T * 1 = 1.  A row of degree d adds d copies of fl(1/d): the rounding errors all
    # have the same sign, so the fp32 sum drifts linearly (d up to 19k here), not like sqrt(d).
    assert float((C1 - 1.0).abs().max()) <= 1e-4
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.rand((n, F), device="cuda", generator=g) * 2 - 1
    y = torch.rand((n, F), device="cuda", generator=g) * 2 - 1
    Cx = mg.spmm(rpt, cit, vat, x, torch.empty((n, F), device="cuda"), plan=plan)
    Cy = mg.spmm(rpt, cit, vat, y, torch.empty((n, F), device="cuda"), plan=plan)
    Cxy = mg.spmm(rpt, cit, vat, x + y, torch.empty((n, F), device="cuda"), plan=plan)
    assert float((Cxy - (Cx + Cy)).abs().max()) <= 1e-5   # linearity (|values| <= 2, rows are convex weights)
    C_noplan = mg.spmm(rpt, cit, vat, x, torch.empty((n, F), device="cuda"))
    assert float((C_noplan - Cx).abs().max()) <= 1e-5
    Cx2 = mg.spmm(rpt, cit, vat, x, torch.empty((n, F), device="cuda"), plan=plan)
    assert torch.equal(Cx, Cx2)                            # deterministic (no atomics)
    # 512 sampled rows (the longest row included) against the fp64 oracle
    deg = np.diff(AT.rowptr.astype(np.int64))
    rows = np.unique(np.concatenate([np.random.default_rng(0).choice(n, 511, replace=False), [int(deg.argmax())]]))
    srp = np.zeros(len(rows) + 1, np.uint64)
    srp[1:] = np.cumsum(deg[rows])
    sci = np.concatenate([AT.colidx[int(AT.rowptr[r]):int(AT.rowptr[r + 1])] for r in rows])
    sva = np.concatenate([AT.vals[int(AT.rowptr[r]):int(AT.rowptr[r + 1])] for r in rows])
    ref = orc.spmm(srp, sci, sva, x.cpu().numpy(), acc64=True)
    got = Cx.cpu().numpy()[rows]
    assert float(np.abs(got - ref).max()) <= 1e-5


@pytest.mark.parametrize("F,late", [(128, False), (128, True), (44, False), (256, True)])
@pytest.mark.parametrize("P,rank", [(4, 1), (8, 7), (2, 0)])
def test_gathered_block_row_matches_oracle(mg, orc, cuda, F, late, P, rank):
    """mggcn_spmm_gathered_f32_i32: one launch over a rank's whole block row with rows stored in stage order and
    global column ids (what the p2p engine runs per aggregation).  Own rows must come from B_own (their copy in
    the gather buffer is poisoned), every other block from the gather buffer; with `late` the stage flags are
    raised by another stream while the kernel is already running, so the kernel has to wait for them."""
    import ctypes as C
    torch = cuda
    rng = np.random.default_rng(5)
    n, deg = 4096, 40
    rp, ci, va = random_csr(rng, n, n, deg, heavy=3)
    cuts = mg.row_cuts(n, P)
    r0, r1 = int(cuts[rank]), int(cuts[rank + 1])
    # block row in stage order: own block first, then blocks rank+1, ... cyclically
    frp = [0]
    fci, fva = [], []
    for r in range(r0, r1):
        c, v = ci[int(rp[r]):int(rp[r + 1])], va[int(rp[r]):int(rp[r + 1])]
        blk = np.searchsorted(cuts, c, side="right") - 1
        stage = (blk - rank) % P
        order = np.argsort(stage, kind="stable")
        fci.append(c[order]); fva.append(v[order])
        frp.append(frp[-1] + len(c))
    frp = np.array(frp, np.uint64)
    fci = np.concatenate(fci).astype(np.uint32); fva = np.concatenate(fva).astype(np.float32)
    ld = (F + 3) // 4 * 4
    B = rng.standard_normal((n, ld)).astype(np.float32)
    want = orc.spmm(rp[r0:r1 + 1] - rp[r0], ci[int(rp[r0]):int(rp[r1])], va[int(rp[r0]):int(rp[r1])], B[:, :F], acc64=True)
    scale = np.abs(orc.spmm(rp[r0:r1 + 1] - rp[r0], ci[int(rp[r0]):int(rp[r1])], np.abs(va[int(rp[r0]):int(rp[r1])]),
                            np.abs(B[:, :F]), acc64=True)).max(axis=1)
    gathered = torch.from_numpy(B).cuda()
    own = gathered[r0:r1].clone()
    gathered[r0:r1] = float("nan")
    rpt, cit, vat = to_dev(torch, frp, fci, fva)
    plan = mg.SpmmPlan(frp.astype(np.uint32), split_nnz=512)
    ws = plan.workspace(F)
    Cm = torch.full((r1 - r0, ld), 7.0, device="cuda")
    flags = torch.zeros(16, dtype=torch.int32, device="cuda") if late else torch.ones(16, dtype=torch.int32, device="cuda")
    err = torch.zeros(4, dtype=torch.int32, device="cuda")
    cuts_h = np.ascontiguousarray(cuts, np.uint32)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    main = torch.cuda.current_stream()
    if late:
        with torch.cuda.stream(side):
            torch.cuda._sleep(20_000_000)   # ~10 ms: the kernel below is resident and waiting by then
            flags.fill_(1)
    mg.check(mg.lib().mggcn_spmm_gathered_f32_i32(
        C.c_void_p(main.cuda_stream), r1 - r0, n, rpt.data_ptr(), cit.data_ptr(), vat.data_ptr(), gathered.data_ptr(),
        own.data_ptr(), ld, Cm.data_ptr(), ld, F, 1.0, 0.0, P, rank, cuts_h.ctypes.data, flags.data_ptr(), err.data_ptr(),
        None, plan.handle, ws.data_ptr() if ws is not None else None))
    torch.cuda.synchronize()
    assert int(err[0]) == 0
    got = Cm.cpu().numpy()[:, :F]
    assert np.isfinite(got).all()
    check(got, want, scale)


@pytest.mark.parametrize("F,ld", [(256, 256), (128, 128), (47, 48), (100, 104), (8, 8)])
@pytest.mark.parametrize("nblk,rot", [(2, 0), (5, 0), (7, 811), (3, 1299)])
def test_column_blocked_plan_matches_oracle(mg, orc, cuda, F, ld, nblk, rot):
    """2-D schedule (mggcn_spmm_plan_create_blocked_i32): feature windows x column blocks over the same CSR arrays,
    C accumulated between blocks, caller's beta on the first block and the epilogue on the last.  Rows carry an
    empty row, hub rows that are split inside a block, and (rot != 0) columns stored in rotated order like the
    engine's grouped operands."""
    torch = cuda
    rng = np.random.default_rng(F + nblk)
    n, m = 1700, 1300
    rp, ci, va = random_csr(rng, n, m, 30, heavy=3)
    if rot:  # store every row ascending in (c - rot) mod m
        for r in range(n):
            a, b = int(rp[r]), int(rp[r + 1])
            o = np.argsort((ci[a:b].astype(np.int64) - rot) % m, kind="stable")
            ci[a:b], va[a:b] = ci[a:b][o], va[a:b][o]
    B = rng.standard_normal((m, ld)).astype(np.float32)
    C0 = rng.standard_normal((n, ld)).astype(np.float32)
    bias = rng.standard_normal(F).astype(np.float32)
    base = orc.spmm(rp, ci, va, B, C_in=C0[:, :F].copy(), alpha=0.5, beta=2.0, acc64=True, F=F)
    want = np.where(base + bias > 0, base + bias, 0.01 * (base + bias))
    absrow = orc.spmm(rp, ci, np.abs(va), np.abs(B), acc64=True, F=F).max(1) * 0.5 + 2 * np.abs(C0[:, :F]).max(1) + 2
    rpt, cit, vat = to_dev(torch, rp, ci, va)
    Bt, bt = torch.from_numpy(B).cuda(), torch.from_numpy(bias).cuda()
    old_b = mg.lib().mggcn_tune(b"blocks", nblk)
    old_w = mg.lib().mggcn_tune(b"window", 32)
    try:
        plan = mg.SpmmPlan(rp.astype(np.uint32), colidx_host=ci, cols=m, F_hint=F, col_rot=rot, col_span=m, split_nnz=256)
        assert plan.blocks == nblk
        Ct = torch.from_numpy(C0).cuda()
        mg.spmm(rpt, cit, vat, Bt, Ct, cols=m, F=F, alpha=0.5, beta=2.0, epi=mg.make_epilogue(mg.EPI_LEAKY_RELU, 0.01, bias=bt),
                plan=plan)
        cfg = mg.lib().mggcn_last_spmm_config().decode()
        assert f"{nblk} column block(s)" in cfg and "spmm_kernel<" in cfg, cfg
        got = Ct.cpu().numpy()
        check(got[:, :F], want, absrow)
        assert np.array_equal(got[:, F:], C0[:, F:])  # padding columns untouched
        # beta = 0 must not read C (NaN-poisoned) and must zero rows that are empty in every block
        Cn = torch.full((n, ld), float("nan"), device="cuda")
        mg.spmm(rpt, cit, vat, Bt, Cn, cols=m, F=F, plan=plan)
        g0 = Cn.cpu().numpy()[:, :F]
        assert np.isfinite(g0).all() and np.all(g0[1] == 0.0)
        check(g0, orc.spmm(rp, ci, va, B, acc64=True, F=F), absrow)
        # deterministic
        Cn2 = torch.empty((n, ld), device="cuda")
        mg.spmm(rpt, cit, vat, Bt, Cn2, cols=m, F=F, plan=plan)
        assert torch.equal(Cn2[:, :F], Cn[:, :F])
    finally:
        mg.lib().mggcn_tune(b"blocks", old_b)
        mg.lib().mggcn_tune(b"window", old_w)


def test_blocked_plan_rejects_unsorted_rows_and_picks_one_block_when_b_fits(mg, cuda):
    rp = np.array([0, 3, 5], np.uint32)
    ci = np.array([4, 2, 7, 0, 1], np.uint32)  # row 0 is not ascending
    old = mg.lib().mggcn_tune(b"blocks", 2)
    try:
        with pytest.raises(mg.MggcnError):
            mg.SpmmPlan(rp, colidx_host=ci, cols=8, F_hint=32)
    finally:
        mg.lib().mggcn_tune(b"blocks", old)
    ci = np.array([2, 4, 7, 0, 1], np.uint32)
    plan = mg.SpmmPlan(rp, colidx_host=ci, cols=8, F_hint=128)  # automatic: a B of 4 KB needs no blocks
    assert plan.blocks == 1


def test_row_pointers_past_2_to_32(mg, orc, cuda):
    """int64 row pointers with VALUES beyond 2^32 (papers100M-shape territory, north_star `matrix.hpp` item): the
    i64 entry point addresses colidx / vals with 64-bit offsets.  The CSR arrays are allocated at their real size
    (2^32 + 12345 + nnz entries: 17 GB each, a tenth of a B200) and the test's entries sit at the far end, so the
    kernel and the plan see genuine row-pointer values beyond 2^32; the entries before them are never read."""
    torch = cuda
    rng = np.random.default_rng(9)
    n, m, F = 900, 700, 128
    rp, ci, va = random_csr(rng, n, m, 35, heavy=2)
    nnz = len(ci)
    off = (1 << 32) + 12345
    big_ci = torch.empty(off + nnz, dtype=torch.int32, device="cuda")
    big_va = torch.empty(off + nnz, dtype=torch.float32, device="cuda")
    big_ci[off:] = torch.from_numpy(ci.view(np.int32)).cuda()
    big_va[off:] = torch.from_numpy(va).cuda()
    rp64 = (rp.astype(np.uint64) + np.uint64(off))
    assert int(rp64[0]) > 2 ** 32
    rpt = torch.from_numpy(rp64.astype(np.int64)).cuda()
    B = rng.standard_normal((m, F)).astype(np.float32)
    Bt = torch.from_numpy(B).cuda()
    want = orc.spmm(rp, ci, va, B, acc64=True)
    absrow = orc.spmm(rp, ci, np.abs(va), np.abs(B), acc64=True).max(1)
    for plan in (None, mg.SpmmPlan(rp64)):
        Ct = torch.empty((n, F), device="cuda")
        mg.spmm(rpt, big_ci, big_va, Bt, Ct, cols=m, F=F, plan=plan)
        check(Ct.cpu().numpy(), want, absrow)
    del big_ci, big_va
    torch.cuda.empty_cache()


@pytest.mark.parametrize("F", [128, 64])
@pytest.mark.parametrize("use_plan", [False, True])
def test_tma_gather_variant_matches_oracle(mg, orc, cuda, F, use_plan):
    """mggcn_tune("tma", 1): rows of B travel through cp.async.bulk into shared memory (experiment of VERDICT r1
    item 6a).  Same contract as the default kernel: alpha/beta, split rows, fused epilogue, empty rows, ragged tails
    of every length modulo the 8-row stage and the 32-entry register chunk."""
    torch = cuda
    rng = np.random.default_rng(F + use_plan)
    n, m = 1100, 900
    rp, ci, va = random_csr(rng, n, m, 37, heavy=3)
    B = rng.standard_normal((m, F)).astype(np.float32)
    C0 = rng.standard_normal((n, F)).astype(np.float32)
    bias = rng.standard_normal(F).astype(np.float32)
    base = orc.spmm(rp, ci, va, B, C_in=C0.copy(), alpha=0.5, beta=2.0, acc64=True)
    want = np.where(base + bias > 0, base + bias, 0.01 * (base + bias))
    absrow = orc.spmm(rp, ci, np.abs(va), np.abs(B), acc64=True).max(1) * 0.5 + 2 * np.abs(C0).max(1) + 2
    rpt, cit, vat = to_dev(torch, rp, ci, va)
    Bt, bt = torch.from_numpy(B).cuda(), torch.from_numpy(bias).cuda()
    plan = mg.SpmmPlan(rp.astype(np.uint32), split_nnz=300) if use_plan else None
    old = mg.lib().mggcn_tune(b"tma", 1)
    try:
        Ct = torch.from_numpy(C0).cuda()
        mg.spmm(rpt, cit, vat, Bt, Ct, cols=m, F=F, alpha=0.5, beta=2.0, epi=mg.make_epilogue(mg.EPI_LEAKY_RELU, 0.01, bias=bt), plan=plan)
        assert "spmm_tma_kernel" in mg.lib().mggcn_last_spmm_config().decode()
        check(Ct.cpu().numpy(), want, absrow)
        Cn = torch.full((n, F), float("nan"), device="cuda")
        mg.spmm(rpt, cit, vat, Bt, Cn, cols=m, F=F, plan=plan)
        got = Cn.cpu().numpy()
        assert np.isfinite(got).all() and np.all(got[1] == 0.0)
        check(got, orc.spmm(rp, ci, va, B, acc64=True), absrow)
    finally:
        mg.lib().mggcn_tune(b"tma", old)
