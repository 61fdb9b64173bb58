"""Dense-transform micro-benchmark on one GPU: the tcgen05 3xTF32 kernel against the FP32 SIMT kernel
of the same library (and torch.matmul with TF32 off = cuBLAS SGEMM, harness only) on the GEMM shapes
of one reddit-shape epoch.  Checks each result against an fp64 product.  One JSON line per case.
  python tools/gemm_bench.py [--n 232968] [--kgroup 1 2 4]"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402


REPS = 10


def timeit(fn, iters=5, warm=2):
    """REPS launches captured into one CUDA graph, so host-side launch cost (ctypes, tensor-map encode)
    is outside the timed region; returns the median time of ONE launch."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(REPS):
            fn()
    g.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); g.replay(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) / REPS)
    return float(np.median(ts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=232968)
    ap.add_argument("--kgroup", type=int, nargs="*", default=[2])
    ap.add_argument("--cublas", action="store_true")
    ap.add_argument("--shapes", default="reddit", choices=["reddit", "products"])
    ap.add_argument("--tune", nargs="*", default=[], help="key=value knobs passed to mggcn_tune")
    ap.add_argument("--presplit", type=int, nargs="*", default=[1], help="weight operand split once per call (1) or per tile (0)")
    ap.add_argument("--split", type=int, nargs="*", default=[0], help="0: both halves rounded to TF32; 1: truncated hi, exact lo")
    args = ap.parse_args()
    mg = load_package()
    L = mg.lib()
    n = args.n
    torch.backends.cuda.matmul.allow_tf32 = False
    # (ta, tb, M, N, K): forward H*W, backward G*W^T, weight gradient X^T*G
    for kv in args.tune:
        k, v = kv.split("=")
        L.mggcn_tune(k.encode(), int(v))
    shapes = [(0, 0, n, 128, 608), (0, 0, n, 128, 128), (0, 0, n, 41, 128), (0, 1, n, 128, 41), (0, 1, n, 128, 128),
              (1, 0, 608, 128, n), (1, 0, 128, 128, n), (1, 0, 128, 41, n)]
    if args.shapes == "products":  # 104 -> 256 -> 256 -> 256 -> 47 (48 padded)
        shapes = [(0, 0, n, 256, 104), (0, 0, n, 256, 256), (0, 0, n, 48, 256), (0, 1, n, 256, 48), (0, 1, n, 256, 256),
                  (1, 0, 104, 256, n), (1, 0, 256, 256, n), (1, 0, 256, 48, n)]
    for ta, tb, M, N, K in shapes:
        g = torch.Generator(device="cuda").manual_seed(M + N + K)
        ldn = (N + 3) // 4 * 4
        A = torch.randn((K, M) if ta else (M, K), device="cuda", generator=g)
        Bfull = torch.randn((N, K) if tb else (K, ldn), device="cuda", generator=g)
        B = Bfull if tb else Bfull[:, :N]
        Cfull = torch.empty((M, ldn), device="cuda")
        Cm = Cfull[:, :N]
        opA = A.t() if ta else A
        opB = B.t() if tb else B
        # fp64 check on a row sample (the full fp64 product of the tall shapes is slow)
        rows = torch.randint(0, M, (min(M, 2048),), device="cuda", generator=g)
        ref = opA[rows].double() @ opB.double()
        scale = (opA[rows].abs().double() @ opB.abs().double()).amax(1, keepdim=True)
        flops = 2.0 * M * N * K
        byts = 4.0 * (M * K + K * N + M * N)
        for tag, tc, kg, ps, sm in [("simt", 0, 0, 1, 0)] + [("tcgen05-3xtf32", 1, k, ps, sm) for k in args.kgroup
                                                            for ps in args.presplit for sm in args.split]:
            L.mggcn_tune(b"gemm_tc", tc)
            L.mggcn_tune(b"gemm_presplit", ps)
            L.mggcn_tune(b"gemm_split", sm)
            if kg:
                L.mggcn_tune(b"gemm_kgroup", kg)
            Cfull.zero_()
            mg.gemm(A, B, Cm, ta=bool(ta), tb=bool(tb))
            torch.cuda.synchronize()
            err = float(((Cm[rows].double() - ref).abs() / scale.clamp_min(1e-30)).max())
            ms = timeit(lambda: mg.gemm(A, B, Cm, ta=bool(ta), tb=bool(tb)))
            print(json.dumps({"tag": tag, "ta": ta, "tb": tb, "M": M, "N": N, "K": K, "kgroup": kg, "presplit": ps, "split": sm, "ms": round(ms, 4),
                              "TFLOPs": round(flops / ms / 1e9, 2), "GBps": round(byts / ms / 1e6, 1),
                              "max_rel_err": err}), flush=True)
        if args.cublas:
            ms = timeit(lambda: torch.matmul(opA, opB))
            print(json.dumps({"tag": "cublas-sgemm(torch)", "ta": ta, "tb": tb, "M": M, "N": N, "K": K, "ms": round(ms, 4),
                              "TFLOPs": round(flops / ms / 1e9, 2)}), flush=True)
    L.mggcn_tune(b"gemm_tc", 1)
    L.mggcn_tune(b"gemm_presplit", 1)
    L.mggcn_tune(b"gemm_split", 0)


if __name__ == "__main__":
    main()
