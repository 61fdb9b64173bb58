"""SpMM micro-benchmark on one GPU: our kernel (knob sweep) vs cuSPARSE (through torch.sparse, harness
only) on a BASELINE.json shape.  Prints one JSON line per measurement.  Usage:
  python tools/spmm_bench.py --shape reddit --F 128 [--sweep] [--blocks P]"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402


def timeit(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts)), float(np.min(ts))


def gathered_bench(mg, AT, args):
    """Kernel cost of the gathered form without any transfer: rank 0's block row (for rank 0 the stage order is
    the column order, so the rows are used as they are), every stage flag already raised."""
    import ctypes as C
    P = args.gathered
    n = AT.n
    cuts = mg.row_cuts(n, P)
    r1 = int(cuts[1])
    e = int(AT.rowptr[r1])
    rp = AT.rowptr[:r1 + 1].astype(np.uint32)
    rpt = torch.from_numpy(rp.view(np.int32)).cuda()
    cit = torch.from_numpy(AT.colidx[:e].view(np.int32)).cuda()
    vat = torch.from_numpy(AT.vals[:e]).cuda()
    plan = mg.SpmmPlan(rp)
    blocks = AT.block_rows(cuts, cuts, only_row=0)
    dev_blocks = []
    for b in blocks:
        brp = b.rowptr.astype(np.uint32)
        dev_blocks.append((torch.from_numpy(brp.view(np.int32)).cuda(), torch.from_numpy(b.colidx.view(np.int32)).cuda(),
                           torch.from_numpy(b.vals).cuda(), mg.SpmmPlan(brp), b.m))
    flags = torch.ones(16, dtype=torch.int32, device="cuda")
    err = torch.zeros(4, dtype=torch.int32, device="cuda")
    cuts_h = np.ascontiguousarray(cuts, np.uint32)
    st = torch.cuda.current_stream().cuda_stream
    print(json.dumps({"gathered_P": P, "rows": r1, "cols": n, "nnz": e}), flush=True)
    for F in args.F:
        ld = (F + 3) // 4 * 4
        B = (torch.rand((n, ld), device="cuda") * 2 - 1)
        own = B[:r1].clone()
        Cm = torch.empty((r1, ld), device="cuda")
        Cs = torch.empty((r1, ld), device="cuda")
        ws = plan.workspace(F)

        def gathered():
            mg.check(mg.lib().mggcn_spmm_gathered_f32_i32(
                C.c_void_p(st), r1, n, rpt.data_ptr(), cit.data_ptr(), vat.data_ptr(), B.data_ptr(), own.data_ptr(), ld,
                Cm.data_ptr(), ld, F, 1.0, 0.0, P, 0, cuts_h.ctypes.data, flags.data_ptr(), err.data_ptr(), None,
                plan.handle, ws.data_ptr() if ws is not None else None))

        def plain():
            mg.spmm(rpt, cit, vat, B, Cm, cols=n, F=F, plan=plan)

        def staged():
            for s, (a, b, c, pl, m) in enumerate(dev_blocks):
                lo = int(cuts[s])
                mg.spmm(a, b, c, B[lo:lo + m], Cs, cols=m, F=F, beta=0.0 if s == 0 else 1.0, plan=pl)
        for tag, fn in (("gathered-1launch", gathered), ("plain-1launch-global-cols", plain),
                        (f"staged-{P}launches", staged)):
            med, mn = timeit(fn)
            print(json.dumps({"tag": tag, "F": F, "ms_med": round(med, 4), "ms_min": round(mn, 4),
                              "gather_GBps": round(e * F * 4 / mn / 1e6, 1)}), flush=True)
        gathered(); staged()
        print(json.dumps({"max_abs_diff_gathered_vs_staged": float((Cm - Cs).abs().max())}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="reddit")
    ap.add_argument("--F", type=int, nargs="+", default=[128])
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--sweep", action="store_true")
    ap.add_argument("--blocks", type=int, default=1, help="P: time the diagonal block (0,0) of a P x P split")
    ap.add_argument("--cusparse", action="store_true")
    ap.add_argument("--splits", type=int, nargs="*", default=[])
    ap.add_argument("--slices", type=int, nargs="*", default=[], help="feature-slice counts: k launches over F/k-wide column windows")
    ap.add_argument("--gathered", type=int, default=0, help="P: time rank 0's block row as ONE gathered launch (flags preset), "
                    "as one plain launch with global columns, and as P staged launches")
    ap.add_argument("--tune", nargs="*", default=[], help="key=value knobs passed to mggcn_tune")
    ap.add_argument("--variants", nargs="*", default=[], help="named knob sets timed one after the other, each "
                    "'tag:key=value,key=value' (knobs are restored afterwards); a 'blocks=' knob rebuilds the plan as a "
                    "column-blocked one, e.g.  base:  tma:tma=1,slices=1  2d:blocks=7,window=32")
    ap.add_argument("--unroll", type=int, default=4)
    ap.add_argument("--threads", type=int, default=64)
    args = ap.parse_args()
    mg = load_package()
    for kv in args.tune:
        k, v = kv.split("=")
        mg.lib().mggcn_tune(k.encode(), int(v))
    t0 = time.time()
    A, _, _, _ = mg.generate(args.shape, seed=1, features=False, scale=args.scale)
    A.normalize(True)
    AT = A.transpose()
    if args.gathered > 1:
        gathered_bench(mg, AT, args)
        return
    if args.blocks > 1:
        cuts = mg.row_cuts(AT.n, args.blocks)
        AT = AT.block_rows(cuts, cuts, only_row=0)[0]
    n, m, nnz = AT.n, AT.m, AT.nnz
    print(json.dumps({"shape": args.shape, "rows": n, "cols": m, "nnz": nnz, "prep_s": round(time.time() - t0, 1)}), flush=True)
    rpt = torch.from_numpy(AT.rowptr.astype(np.uint32).view(np.int32)).cuda()
    cit = torch.from_numpy(AT.colidx.view(np.int32)).cuda()
    vat = torch.from_numpy(AT.vals).cuda()
    plan = mg.SpmmPlan(AT.rowptr.astype(np.uint32))
    for F in args.F:
        ld = (F + 3) // 4 * 4
        B = (torch.rand((m, ld), device="cuda") * 2 - 1)
        Cm = torch.empty((n, ld), device="cuda")
        b_alg = nnz * 8 + (n + 1) * 4 + m * F * 4 + n * F * 4
        b_gather = nnz * F * 4 + nnz * 8 + n * F * 4

        def report(tag, med, mn, **kw):
            print(json.dumps({"tag": tag, "F": F, "ms_med": round(med, 4), "ms_min": round(mn, 4),
                              "alg_GBps": round(b_alg / mn / 1e6, 1), "gather_GBps": round(b_gather / mn / 1e6, 1),
                              "GFLOPs": round(2 * nnz * F / mn / 1e6, 1), **kw}), flush=True)

        configs = [(args.unroll, args.threads, 1, True)]
        if args.sweep:
            configs = [(u, t, p, pl) for u in (4, 8) for t in (64, 128, 256) for p in (0, 1) for pl in (True,)] + [(8, 128, 1, False)]
        for unroll, threads, policy, use_plan in configs:
            mg.lib().mggcn_tune(b"unroll", unroll)
            mg.lib().mggcn_tune(b"threads", threads)
            mg.lib().mggcn_tune(b"policy", policy)
            med, mn = timeit(lambda: mg.spmm(rpt, cit, vat, B, Cm, cols=m, F=F, plan=plan if use_plan else None))
            report("ours", med, mn, unroll=unroll, threads=threads, policy=policy, plan=use_plan)
        for sp in args.splits:
            pl = mg.SpmmPlan(AT.rowptr.astype(np.uint32), split_nnz=sp)
            med, mn = timeit(lambda: mg.spmm(rpt, cit, vat, B, Cm, cols=m, F=F, plan=pl))
            report("ours-split", med, mn, split=sp)
        for k in args.slices:
            if F % (4 * k):
                continue
            w = F // k

            def sliced():
                for i in range(k):
                    mg.spmm(rpt, cit, vat, B[:, i * w:(i + 1) * w], Cm[:, i * w:(i + 1) * w], cols=m, F=w, plan=plan)
            med, mn = timeit(sliced)
            report("ours-sliced", med, mn, slices=k)
        for spec in args.variants:
            tag, _, kvs = spec.partition(":")
            knobs = dict(kv.split("=") for kv in kvs.split(",") if kv)
            saved = {k: mg.lib().mggcn_tune(k.encode(), int(v)) for k, v in knobs.items()}
            try:
                pl = plan
                if "blocks" in knobs:
                    pl = mg.SpmmPlan(AT.rowptr.astype(np.uint32), colidx_host=AT.colidx, cols=m, F_hint=F)
                fn = lambda: mg.spmm(rpt, cit, vat, B, Cm, cols=m, F=F, plan=pl)
                fn()
                cfg = mg.lib().mggcn_last_spmm_config().decode()
                med, mn = timeit(fn)
                report("variant-" + tag, med, mn, knobs=knobs, config=cfg, plan_blocks=pl.blocks)
                ref_plain = mg.spmm(rpt, cit, vat, B, torch.empty_like(Cm), cols=m, F=F, plan=None)
                print(json.dumps({"tag": "variant-" + tag, "max_abs_diff_vs_plain": float((ref_plain[:, :F] - Cm[:, :F]).abs().max())}), flush=True)
            finally:
                for k, v in saved.items():
                    mg.lib().mggcn_tune(k.encode(), v)
        if args.cusparse:
            S = torch.sparse_csr_tensor(rpt.to(torch.int32), cit, vat, size=(n, m))
            Bc = B[:, :F].contiguous()
            med, mn = timeit(lambda: torch.sparse.mm(S, Bc))
            report("cusparse(torch.sparse.mm)", med, mn)
            ref = torch.sparse.mm(S, Bc)
            mg.spmm(rpt, cit, vat, B, Cm, cols=m, F=F, plan=plan)
            print(json.dumps({"max_abs_diff_vs_cusparse": float((ref - Cm[:, :F]).abs().max())}), flush=True)


if __name__ == "__main__":
    main()
