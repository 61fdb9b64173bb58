#!/usr/bin/env python
"""bench.py -- GCN epoch time on the reddit-shape synthetic graph (BASELINE.json configs[1]).

  python bench.py --gpus N --steps K --warmup W              our arm (one process per GPU; torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K ...    the reference arm (rank 0 only)

A "step" is one epoch: forward (4 layers: GEMM + staged SpMM each), loss, backward, Adam -- the body
of the reference's epoch loop (src/main.cpp:158-169).  `value` is ms per epoch with X, Y, the
adjacency blocks and the weights resident in HBM; `e2e` is the same epoch entered through the C-ABI
with X and Y copied from pinned host memory every step and the loss read back.

Rank 0 prints ONE JSON line.  torch is used for the process group, pinned memory and events only.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402

ADAM = dict(lr=1e-2, beta1=0.9, beta2=0.999, wd=5e-4, eps=1e-8)  # src/main.cpp:126,163

_JSON_FD = None


def claim_stdout():
    """Libraries print banners on fd 1 (NCCL's version line, torchrun's notes): keep the real stdout for the ONE
    JSON line and send everything else to stderr."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(line.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, line)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--shape", default="reddit")
    ap.add_argument("--hidden", type=int, nargs="+", default=[128, 128, 128])
    ap.add_argument("--comm", default="p2p-grouped", choices=["nccl", "p2p", "p2p-staged", "p2p-grouped"],
                    help="exchange back-end at N > 1: p2p-grouped (measured fastest: 3.27 ms at 8 GPUs), p2p-staged (3.54), "
                         "p2p = gathered single launch (3.75), nccl")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-overlap", action="store_true")
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the graph (debugging only; marks the line)")
    ap.add_argument("--cpu-sample", type=float, default=1.0, help="row fraction timed on the host cores")
    ap.add_argument("--cpu-repeats", type=int, default=3)
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--tune", nargs="*", default=[], help="key=value kernel knobs (mggcn_tune), experiments only")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm = [float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 8:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def spmm_alg_bytes(rows, cols, nnz, F, beta_nonzero=False, w=4):
    """every operand touched exactly once."""
    return nnz * 8 + (rows + 1) * w + cols * F * 4 + rows * F * 4 * (2 if beta_nonzero else 1)


def layer_plan(sizes):
    """The SpMM / GEMM calls of one epoch in the reference's order (src/gcn.hpp:556-608), as
    (kind, F or (in,out)) tuples -- used for byte / flop accounting and the host-core baseline."""
    L = len(sizes) - 1
    ops = []
    for l in range(L):
        i, o = sizes[l], sizes[l + 1]
        ops += [("gemm", (i, o)), ("spmm", o)] if o <= i else [("spmm", i), ("gemm", (i, o))]
    for l in range(L - 1, -1, -1):
        i, o = sizes[l], sizes[l + 1]
        if o <= i:
            if l != 0:
                ops.append(("spmm", o))
            ops.append(("gemm_tn", (i, o)))
            if l != 0:
                ops.append(("gemm", (o, i)))
        else:
            ops.append(("gemm_tn", (i, o)))
            if l != 0:
                ops += [("gemm", (o, i)), ("spmm", i)]
    return ops


def cpu_epoch_sample(mg, A, A_T, X, sizes, frac, repeats=1):
    """Host-core baseline ("port"): the oracle's kernels on the first frac*N output rows of every
    SpMM and GEMM of one epoch (all columns / full B), scaled by 1/frac.  Every op of the epoch is
    row-parallel and the vertex order is a random permutation, so cost is linear in rows."""
    from oracle import oracle as orc
    N = A.n
    ns = max(8, int(N * frac))
    rng = np.random.default_rng(0)
    subs = []
    for M in (A_T, A):
        e = int(M.rowptr[ns])
        subs.append((M.rowptr[:ns + 1], M.colidx[:e], M.vals[:e]))
    t_spmm = t_gemm = 0.0
    widths = {}
    n_fwd_spmm = len(sizes) - 1
    k = 0
    for kind, arg in layer_plan(sizes) * repeats:
        k %= 2 * n_fwd_spmm - 1
        if kind == "spmm":
            F = arg
            B = widths.setdefault(F, rng.standard_normal((N, F)).astype(np.float32))
            rp, ci, va = subs[0] if k < n_fwd_spmm else subs[1]
            k += 1
            t0 = time.perf_counter()
            orc.spmm(rp, ci, va, B)
            t_spmm += time.perf_counter() - t0
        elif kind == "gemm":
            i, o = arg
            a = rng.standard_normal((ns, i)).astype(np.float32)
            w = rng.standard_normal((i, o)).astype(np.float32)
            t0 = time.perf_counter()
            orc.gemm(a, w)
            t_gemm += time.perf_counter() - t0
        else:
            i, o = arg
            a = rng.standard_normal((ns, i)).astype(np.float32)
            g = rng.standard_normal((ns, o)).astype(np.float32)
            t0 = time.perf_counter()
            orc.gemm(a, g, ta=True)
            t_gemm += time.perf_counter() - t0
    scale = N / ns / repeats
    return {"value": (t_spmm + t_gemm) * scale * 1e3, "unit": "ms/epoch", "cores": orc.threads(), "kind": "port",
            "sample": f"oracle SpMM+GEMM kernels of one epoch on the first {ns} of {N} output rows, {repeats} pass(es) (x{scale:.2f}); "
                      f"SpMM share {t_spmm / max(t_spmm + t_gemm, 1e-9):.2f}; measured {t_spmm + t_gemm:.1f} s"}


# ------------------------------------------------------------------------------------------------
def run_reference(args, mg, rank):
    """The reference arm: the UNMODIFIED reference program (oracle/_ref/mg_gcn_ref, built from
    /root/reference/src where it lies) through its own CLI on the same generated dataset.  Its only
    implementation is CUDA (cuSPARSE/cuBLAS/NCCL); there is no CPU path in the reference, so when the
    binary or a GPU is missing the arm falls back to the oracle port on the host cores."""
    if rank != 0:
        return
    P = args.gpus
    A, X, Y, C = mg.generate(args.shape, seed=1, scale=args.scale)
    sizes = [X.shape[1]] + args.hidden + [mg.pad_classes(C, P)]
    exe = os.path.join(ROOT, "oracle", "_ref", "mg_gcn_ref")
    base = {"metric": "gcn_epoch_ms", "unit": "ms/epoch", "n_gpus": P, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "impl": "reference",
            "config": {"workload": f"{args.shape}-shape synthetic, N={A.n}, nnz={A.nnz}, F={X.shape[1]}, "
                                   f"{len(args.hidden)}x{args.hidden[0]} hidden, {C} classes", "P": P,
                       "l2": "inputs larger than L2 (CSR 0.9 GB per operand)"}}
    have_gpu = False
    try:
        import torch
        have_gpu = torch.cuda.is_available() and torch.cuda.device_count() >= P
    except Exception:
        pass
    if os.path.exists(exe) and have_gpu:
        with tempfile.TemporaryDirectory(prefix="mggcn_ref_") as d:
            ds = os.path.join(d, "permuted", args.shape)
            mg.write_dataset(ds, A, X, Y)
            os.makedirs(os.path.join(d, "csvs"))
            E = args.warmup + args.steps
            cmd = [exe, "-P", str(P), "-R", "1", "-E", str(E), "train", ds + "/", str(len(args.hidden))] + [str(h) for h in args.hidden]
            t0 = time.time()
            r = subprocess.run(cmd, cwd=d, capture_output=True, text=True, timeout=3000)
            wall = time.time() - t0
            ep = []
            for line in r.stderr.splitlines():
                f = line.split()
                if len(f) == 4 and f[0].isdigit():
                    ep.append((int(f[0]), float(f[1]), float(f[2]), float(f[3])))
            if r.returncode == 0 and len(ep) == E:
                times = [t for (_, _, _, t) in ep[args.warmup:]]
                ms = float(np.mean(times)) * 1e3
                base.update({"value": ms, "ms_per_step": ms,
                             "e2e": {"value": ms, "unit": "ms/epoch", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                             "cpu_baseline": {"value": ms, "unit": "ms/epoch", "cores": 1, "kind": "reference",
                                              "sample": "the reference's own CUDA build (cuSPARSE/cuBLAS/NCCL; it has no CPU path), "
                                                        f"all {E} epochs, mean of the last {args.steps}; one host thread drives {P} GPU(s)"},
                             "reference_detail": {"cmd": " ".join(cmd[1:]), "median_ms": float(np.median(times)) * 1e3,
                                                  "min_ms": float(np.min(times)) * 1e3, "losses": [l for (_, l, _, _) in ep][:12],
                                                  "wall_s": wall}})
                emit(base)
                return
            base["reference_binary_failed"] = (r.stderr or r.stdout)[-400:]
    # fallback: oracle port on the host cores
    A.normalize(True)
    A_T = A.transpose()
    cb = cpu_epoch_sample(mg, A, A_T, X, sizes, args.cpu_sample, args.cpu_repeats)
    base.update({"value": cb["value"], "ms_per_step": cb["value"], "cpu_baseline": cb,
                 "e2e": {"value": cb["value"], "unit": "ms/epoch", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    emit(base)


# ------------------------------------------------------------------------------------------------
def main():
    args = parse()
    claim_stdout()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:  # every rank pre-processes the graph: share the host cores instead of oversubscribing them
        if os.environ.get("OMP_NUM_THREADS", "1") == "1":  # torchrun's default of 1 makes the host prep 8x slower
            os.environ["OMP_NUM_THREADS"] = str(max(1, (os.cpu_count() or 8) // world))
    mg = load_package()
    if args.impl == "reference":
        run_reference(args, mg, rank)
        return
    for kv in args.tune:
        k, v = kv.split("=")
        mg.lib().mggcn_tune(k.encode(), int(v))
    import torch
    import torch.distributed as dist
    assert torch.cuda.is_available(), "bench.py needs a GPU: the product has no CPU fallback"
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    P = world

    # ---- dataset and host pre-processing (src/main.cpp:82-85,139-149); every rank builds its block row
    t0 = time.time()
    A, X, Y, C = mg.generate(args.shape, seed=1, scale=args.scale)
    N, nnz = A.n, A.nnz
    sizes = [X.shape[1]] + args.hidden + [mg.pad_classes(C, P)]
    nccl_id = None
    comm = {"p2p": mg.COMM_P2P, "p2p-staged": mg.COMM_P2P_STAGED, "p2p-grouped": mg.COMM_P2P_GROUPED,
            "nccl": mg.COMM_NCCL}[args.comm]
    if P > 1 and comm == mg.COMM_NCCL:
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            buf = (__import__("ctypes").c_char * 128)()
            mg.check(mg.lib().mggcn_nccl_unique_id(buf))
            idt.copy_(torch.frombuffer(bytearray(bytes(buf)), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        nccl_id = bytes(idt.cpu().numpy().tobytes())
    eng, A_T = mg.build_rank_engine(A, sizes, world=P, rank=rank, device=local, use_graph=not args.no_graph,
                                    nccl_id=nccl_id, comm=comm, overlap=not args.no_overlap)
    if P > 1 and comm != mg.COMM_NCCL:
        h = eng.comm_handle()
        ht = torch.frombuffer(bytearray(h), dtype=torch.uint8).cuda()
        allh = [torch.empty_like(ht) for _ in range(P)]
        dist.all_gather(allh, ht)
        eng.comm_connect([bytes(t.cpu().numpy().tobytes()) for t in allh])
    cuts = mg.row_cuts(N, P)
    r0, r1 = int(cuts[rank]), int(cuts[rank + 1])
    Xl = torch.from_numpy(np.ascontiguousarray(X[r0:r1])).pin_memory()
    Yl = torch.from_numpy(np.ascontiguousarray(Y[r0:r1])).pin_memory()
    eng.set_inputs(Xl, Yl)
    prep_s = time.time() - t0
    sc = torch.cuda.ExternalStream(eng.stream(0))

    def barrier():
        eng.sync()
        torch.cuda.synchronize()
        if P > 1:
            dist.barrier()

    def timed(fn, steps):
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.perf_counter()
        a.record(sc)
        for _ in range(steps):
            fn()
        b.record(sc)
        eng.sync()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - w0) * 1e3
        t = torch.tensor([a.elapsed_time(b), wall], device="cuda", dtype=torch.float64)
        if P > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]) / steps, float(t[1]) / steps

    # ---- resident-input epoch (value)
    for _ in range(max(args.warmup, 3)):
        eng.epoch(**ADAM)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = mg.lib().mggcn_launch_count()
    ms_dev, ms_wall = timed(lambda: eng.epoch(**ADAM), args.steps)
    launches = mg.lib().mggcn_launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    loss_sum, correct = eng.loss()

    # ---- end to end through the C-ABI: host X, Y in, loss out, every step
    # (every step copies X and Y from pinned host memory: the copy of step i+1's inputs is issued as a
    # prefetch before epoch i is launched, so it travels over PCIe while epoch i computes, and is swapped
    # in on the device before epoch i+1 starts; K steps in the timed region = K copies, K epochs, K reads)
    def e2e_step():
        eng.epoch(**ADAM)                  # swaps in the inputs prefetched during the previous step
        eng.set_inputs(Xl, Yl, async_=2)   # next step's inputs, in flight while this epoch computes
        eng.loss()                         # device -> host read of this step's loss (synchronises)
    eng.set_inputs(Xl, Yl, async_=2)
    e2e_step()
    e2e_dev, e2e_wall = timed(e2e_step, args.steps)

    # ---- dominant kernel, live: eager pass with the engine's named timers (CUDA events on its stream)
    eng.set_mode(use_graph=False, timers=True)
    per = {}
    for _ in range(3):
        eng.epoch(**ADAM)
        eng.sync()
        for line in eng.timers().splitlines():
            name, ms = line.rsplit(":", 1)
            per.setdefault(name, []).append(float(ms))
    eng.set_mode(use_graph=not args.no_graph, timers=False)
    med = {k: float(np.median(v)) for k, v in per.items()}
    # forward SpMMs of width 128 through the forward operand (layers whose SpMM width is hidden[0])
    n_loc = r1 - r0
    F = args.hidden[0]
    fused = P > 1 and args.comm == "p2p"
    dom_nnz = A_T.nnz / (P * P if (P > 1 and not fused) else P)
    if P == 1 or fused:
        dom = [v for k, v in med.items() if k.endswith("_0_matmul-spmm") and not k.startswith(f"{len(sizes) - 2}_")]
        dom_ms = float(np.mean(dom)) if dom else float("nan")
        alg = spmm_alg_bytes(n_loc, N, dom_nnz, F)
        sliced = 90e6 < N * F * 4 <= 124e6 and F % 8 == 0 and F >= 128 and not any(t.startswith("slices=") for t in args.tune)
        dom_name = ((f"spmm_kernel<16,1,4> x2 (two {F // 2}-wide column windows, automatic feature slicing) forward, F={F}, whole matrix"
                     if sliced else f"spmm_kernel<32,1,4> forward, F={F}, whole matrix") if P == 1 else
                    f"spmm_kernel<32,1,4> own block + spmm_kernel<32,1,4,gathered> remote blocks, forward, F={F}, block row {n_loc}x{N} "
                    "(the gathered launch waits on-device for the peers' blocks)")
    else:
        grouped = args.comm == "p2p-grouped"  # launches differ in size there: take the own-block launch (stage 0)
        dom = [v for k, v in med.items() if k.endswith("_matmul-spmm-stage") and k.split("_")[1] == "0"
               and not k.startswith(f"{len(sizes) - 2}_") and (not grouped or k.split("_")[2] == "0")]
        dom_ms = float(np.mean(dom)) if dom else float("nan")
        blk_nnz = A_T.nnz / (P * P)
        alg = blk_nnz * 8 + (n_loc + 1) * 4 + n_loc * F * 4 + n_loc * F * 4 * 2
        dom_name = f"spmm_kernel<32,1,4> forward stage, F={F}, one {n_loc}x{n_loc} block"
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "spmm_traffic.json"))).get(f"P{P}_F{F}")
    except (OSError, ValueError):
        pass
    achieved = alg / (dom_ms * 1e-3) / 1e9 if dom_ms == dom_ms and dom_ms > 0 else None
    spmm_ms = sum(v for k, v in med.items() if k.endswith("matmul-spmm"))
    gemm_ms = sum(v for k, v in med.items() if k.endswith("matmul-gemm"))
    xfer_ms = sum(v for k, v in med.items() if k.endswith("matmul-bcast"))

    if rank != 0:
        eng.close()
        if P > 1:
            dist.destroy_process_group()
        return

    cpu = None
    if P == 1 and not args.skip_cpu:
        A_full = A  # normalised in place by build_rank_engine
        cpu = cpu_epoch_sample(mg, A_full, A_T, X, sizes, args.cpu_sample, args.cpu_repeats)

    out = {
        "metric": "gcn_epoch_ms", "value": ms_dev, "unit": "ms/epoch", "n_gpus": P, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_dev, "higher_is_better": False, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.shape}-shape synthetic, N={N}, nnz={nnz}, F={X.shape[1]}, "
                               f"{len(args.hidden)}x{args.hidden[0]} hidden, {C} classes (padded to {sizes[-1]})",
                   "P": P, "partition": "1-D row blocks, replicated weights", "comm": args.comm if P > 1 else "none",
                   "graph": not args.no_graph, "l2": "inputs larger than L2 (CSR 0.9 GB per operand, X 0.57 GB)",
                   "scale": args.scale, **({"tune": args.tune} if args.tune else {})},
        "e2e": {"value": e2e_wall, "unit": "ms/epoch", "device_ms": e2e_dev,
                "h2d_bytes_per_step": int(N * X.shape[1] * 4 + N * 4), "d2h_bytes_per_step": 16 * P},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": dom_name, "achieved": achieved, "peak": peak,
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s (B200_PROFILING.md)",
                     "unit": "GB/s", "frac": (achieved / peak) if achieved else None, "traffic": traffic,
                     "alg_bytes_per_launch": alg, "launch_ms": dom_ms,
                     "gather_GBps": (dom_nnz * F * 4 / (dom_ms * 1e-3) / 1e9) if dom_ms > 0 else None,
                     "GFLOPs": (2 * dom_nnz * F / (dom_ms * 1e-3) / 1e9) if dom_ms > 0 else None,
                     "note": "B (119 MB at F=128) is gathered avg-degree (~490) times: the kernel is bound by the L2->SM "
                             "return path (l1tex 88 %, lts 84 % of peak in ncu), not by HBM (SURVEY.md hard part 1); "
                             "gather_GBps is the operand traffic it sustains; achieved/frac use the algorithmic bytes "
                             "of ONE pass over the operands and launch_ms is the whole aggregation"},
        "breakdown_ms": {"spmm": spmm_ms, "gemm": gemm_ms, "exchange": xfer_ms,
                         "loss": sum(v for k, v in med.items() if k.endswith("loss-layer")),
                         "adam": med.get("adam-update"), "allreduce": med.get("allreduce"),
                         "eager_epoch_sum": sum(v for k, v in med.items() if not k.endswith(("-stage", "-bcast")))},
        "wall_ms_per_step": ms_wall, "prep_s": prep_s,
        "loss": loss_sum / N if P == 1 else None,
    }
    if P > 1:
        # exchange against the NVLink roofline (BASELINE.json north_star): bytes this rank receives per epoch =
        # sum over the aggregations of (P-1)/P * N * ld * 4, over the time its copies are in flight (eager timers)
        try:
            widths = [a for k, a in layer_plan(sizes) if k == "spmm"]
            rx = sum((P - 1) / P * N * ((w + 3) // 4 * 4) * 4 for w in widths)
            out["exchange"] = {"rx_bytes_per_rank_per_epoch": int(rx), "copy_ms_per_epoch": xfer_ms,
                               "GBps": (rx / (xfer_ms * 1e-3) / 1e9) if xfer_ms > 0 else None,
                               "nvlink_peak_GBps": 900.0,
                               "nvlink_frac": (rx / (xfer_ms * 1e-3) / 1e9 / 900.0) if xfer_ms > 0 else None,
                               "note": "copies run on the exchange stream behind the SpMM launches of earlier stages; "
                                       "spmm in breakdown_ms already contains any time a launch waited for a block"}
        except Exception as ex:  # accounting only: never lose the line
            out["exchange"] = {"error": repr(ex)}
    if cpu:
        out["cpu_baseline"] = cpu
    emit(out)
    eng.close()
    if P > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
