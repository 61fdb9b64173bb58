#!/usr/bin/env python
"""bench.py -- GCN epoch time on the synthetic graphs of BASELINE.json (default: configs[1], reddit-shape 3x128).

  python bench.py --gpus N --steps K --warmup W [--shape reddit|products|proteins|papers100M]
        our arm (one process per GPU; torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K ...
        the reference arm (rank 0 only): the unmodified reference program through its own CLI

A "step" is one epoch: forward (GEMM + staged SpMM per layer), loss, backward, Adam -- the body of the
reference's epoch loop (src/main.cpp:158-169).  `value` is ms per epoch with X, Y, the adjacency blocks and the
weights resident in HBM; `e2e` is the same epoch entered through the C-ABI with X and Y copied from pinned host
memory every step and the loss read back.  Before anything is timed the engine trains 10 epochs from the
reference's initialisation and the line carries that loss curve next to the reference binary's at the same P
(`parity`), so a fast number for a wrong computation cannot pass unnoticed.

Rank 0 prints ONE JSON line.  torch is used for the process group, pinned memory and events only.
"""
import argparse
import hashlib
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
PARITY_EPOCHS = 10
SEED = 1

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


def log(*a):
    print("[bench]", *a, file=sys.stderr, flush=True)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--shape", default="reddit", choices=["reddit", "products", "proteins", "papers100M", "oracle4k"])
    ap.add_argument("--hidden", type=int, nargs="+", default=None, help="default: the config's own (3x128, products 3x256)")
    ap.add_argument("--comm", default="p2p-grouped", choices=["nccl", "p2p", "p2p-staged", "p2p-grouped"],
                    help="exchange back-end at N > 1: p2p-grouped (measured fastest), p2p-staged, "
                         "p2p = gathered single launch, nccl = the reference's broadcast schedule")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-overlap", action="store_true")
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the graph (debugging only; marks the line)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="host-core baseline: CPU work to aim for")
    ap.add_argument("--cpu-sample", type=float, default=0.0, help="row fraction timed on the host cores (0 = from --cpu-seconds)")
    ap.add_argument("--cpu-repeats", type=int, default=1)
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--skip-parity", action="store_true", help="no 10-epoch loss curve / reference run (experiments only)")
    ap.add_argument("--e2e-variants", action="store_true", help="also time the end-to-end step with the SM-driven input copy (always on at N > 1)")
    ap.add_argument("--e2e-sweep", nargs="*", default=[], help="knob combinations 'k=v,k=v' timed resident and end to end (experiments)")
    ap.add_argument("--skip-exposed", action="store_true", help="no exposed-communication measurement at N > 1")
    ap.add_argument("--tune", nargs="*", default=[], help="key=value kernel knobs (mggcn_tune), experiments only")
    a = ap.parse_args()
    return a


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
        pw = [float(r[3]) for r in self.rows if len(r) >= 8 and r[3].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 8:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w": float(np.median(pw)) if pw else None, "reasons": sorted(reasons), "samples": len(sm)}


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


def workload_text(shape, N, nnz, F, hidden, C, scale=1.0):
    """The same string in both arms (the driver compares `config` textually).  Where the reference arm cannot
    build the graph (papers100M-shape) both arms print the generator's nominal entry count."""
    if shape in REF_UNSUPPORTED and scale == 1.0:
        nnz = f"{SHAPE_NOMINAL_NNZ[shape]} (nominal)"
    return (f"{shape}-shape synthetic, N={N}, nnz={nnz}, F={F}, {len(hidden)}x{hidden[0]} hidden, {C} classes; "
            "inputs larger than L2")


def hidden_of(args, mg):
    return list(args.hidden) if args.hidden else list(mg.HIDDEN[args.shape])


def cpu_epoch_sample(fwd, bwd, sizes, frac, repeats=1, seconds=15.0, dense_rows=None):
    """Host-core baseline ("port"): the oracle's kernels on the first frac*N output rows of every
    SpMM and GEMM of one epoch (all columns / full B), scaled by 1/frac.  Every op of the epoch is
    row-parallel and the vertex order is a random permutation, so cost is linear in rows.
    fwd / bwd: the two aggregation operands as host CSR (rows of A_T and of A).  frac = 0: a 0.5 % probe first,
    then the fraction that should take about `seconds` of CPU work.  dense_rows < columns: B is kept that many rows
    high and column ids wrap around it (papers100M-shape: a full-height B is 57 GB; 4 GB is still far beyond any
    cache, so the gathers stay DRAM-bound random row reads)."""
    from oracle import oracle as orc
    N = fwd.n
    M = fwd.m if dense_rows is None else min(dense_rows, fwd.m)  # rows of the dense operand B
    rng = np.random.default_rng(0)

    def run(ns, reps):
        subs = []
        for Mx in (fwd, bwd):
            e = int(Mx.rowptr[ns])
            ci = Mx.colidx[:e] if M == Mx.m else (Mx.colidx[:e] % np.uint32(M)).astype(np.uint32)
            subs.append((Mx.rowptr[:ns + 1], ci, Mx.vals[:e]))
        t_spmm = t_gemm = 0.0
        widths = {}
        n_fwd = len(sizes) - 1
        k = 0
        for kind, arg in layer_plan(sizes) * reps:
            k %= 2 * n_fwd - 1
            if kind == "spmm":
                F = arg
                B = widths.get(F)
                if B is None:
                    B = widths[F] = rng.standard_normal((M, F), dtype=np.float32)
                rp, ci, va = subs[0] if k < n_fwd else subs[1]
                k += 1
                t0 = time.perf_counter()
                orc.spmm(rp, ci, va, B)
                t_spmm += time.perf_counter() - t0
            else:
                i, o = arg
                a = rng.standard_normal((ns, i), dtype=np.float32)
                b = rng.standard_normal((i, o) if kind == "gemm" else (ns, o), dtype=np.float32)
                t0 = time.perf_counter()
                orc.gemm(a, b, ta=(kind != "gemm"))
                t_gemm += time.perf_counter() - t0
        return t_spmm, t_gemm

    if frac <= 0:
        probe = max(8, int(N * 0.005))
        ts, tg = run(probe, 1)
        full = (ts + tg) * N / probe
        frac = min(1.0, max(0.005, seconds / max(full, 1e-9)))
        repeats = max(1, int(seconds / max(full, 1e-9))) if frac >= 1.0 else 1
        repeats = min(repeats, 5)
    ns = max(8, min(N, int(N * frac)))
    t_spmm, t_gemm = run(ns, repeats)
    scale = N / ns / repeats
    return {"value": (t_spmm + t_gemm) * scale * 1e3, "unit": "ms/epoch", "cores": orc.threads(), "kind": "port",
            "sample": f"oracle SpMM+GEMM kernels of one epoch on the first {ns} of {N} output rows, {repeats} pass(es) (x{scale:.2f}); "
                      f"SpMM share {t_spmm / max(t_spmm + t_gemm, 1e-9):.2f}; measured {t_spmm + t_gemm:.1f} s"}


# ------------------------------------------------------------------------------------------------
# reference binary: timed arm and loss-curve checker
# ------------------------------------------------------------------------------------------------
REF_EXE = os.path.join(ROOT, "oracle", "_ref", "mg_gcn_ref")
# stand-alone dataset writer (generator + file objects linked in, no libmggcn.so): the reference arm generates its
# files with it, so that nothing of this repository's library is mapped into the process that times the reference
GEN_EXE = os.path.join(ROOT, "mg-gcn_b200", "cpp", "bin", "mggcn_gen")


def maps_libmggcn():
    """Is this repository's library mapped into the calling process?  (The reference arm reports it: it must not be.)"""
    try:
        return any("libmggcn" in line for line in open("/proc/self/maps"))
    except OSError:
        return None


def write_dataset_standalone(mg, shape, scale, ds):
    """Runs mggcn_gen for `shape` into directory `ds`; returns {"n", "nnz", "F", "classes"} or None.
    Same generator code and arguments as mg.generate + mg.write_dataset: the files are byte-identical
    (tests/test_bench_contract.py)."""
    if not os.path.exists(GEN_EXE):
        return None
    n_raw, edges, f_raw, classes, model, gamma, max_deg = mg.shape_params(shape, scale)
    cmd = [GEN_EXE, str(n_raw), str(edges), str(f_raw), str(classes), str(model), repr(float(gamma)), str(max_deg), str(SEED),
           str(int(shape not in mg.UNPERMUTED)), ds]
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=3000)
        return json.loads(r.stdout.strip().splitlines()[-1]) if r.returncode == 0 else None
    except (subprocess.TimeoutExpired, ValueError, IndexError, OSError):
        return None
REF_UNSUPPORTED = {
    "papers100M": "the reference cannot run papers100M-shape: 3.23e9 entries exceed its signed 32-bit cuSPARSE indices "
                  "(CUSPARSE_INDEX_32I from x_t = v_t = unsigned, src/main.cpp:43-45, src/mg_gcn.hpp:93-100) and its "
                  "32-bit PIGO ordinal (src/matrix.hpp:225); N*F = 1.4e10 also overflows the 32-bit products of its "
                  "dense containers; nothing below 2^31 entries is left once the graph is loaded whole on one host",
}


SHAPE_NOMINAL_NNZ = {"papers100M": 3231371744}


def curve_cache_path(shape, scale, P, hidden):
    key = hashlib.sha1(json.dumps([shape, scale, P, hidden, SEED]).encode()).hexdigest()[:16]
    d = os.path.join(tempfile.gettempdir(), "mggcn_parity")
    os.makedirs(d, exist_ok=True)
    return os.path.join(d, f"refcurve_{shape}_P{P}_{key}.json")


def have_gpus(P):
    try:
        import torch
        return torch.cuda.is_available() and torch.cuda.device_count() >= P
    except Exception:
        return False


def run_reference_binary(mg, shape, scale, P, hidden, epochs, dataset=None):
    """Runs oracle/_ref/mg_gcn_ref (the reference's own sources, compiled where they lie) through its CLI on the
    generated dataset.  Returns (epoch tuples [(e, loss, acc, secs)], info dict) or (None, {"why": ...})."""
    if shape in REF_UNSUPPORTED and scale == 1.0:
        return None, {"why": REF_UNSUPPORTED[shape]}
    if not os.path.exists(REF_EXE):
        return None, {"why": "oracle/_ref/mg_gcn_ref is not built (needs /root/reference at build time)"}
    if not have_gpus(P):
        return None, {"why": f"the reference is a CUDA program and needs {P} GPU(s)"}
    with tempfile.TemporaryDirectory(prefix="mggcn_ref_") as d:
        ds = os.path.join(d, "permuted", shape)
        t0 = time.time()
        os.makedirs(os.path.dirname(ds), exist_ok=True)
        meta = write_dataset_standalone(mg, shape, scale, ds) if dataset is None else None
        standalone = meta is not None
        if meta is None:
            if dataset is None:
                dataset = mg.generate(shape, seed=SEED, scale=scale, permute=shape not in mg.UNPERMUTED)
            A, X, Y, C = dataset
            mg.write_dataset(ds, A, X, Y)
            meta = {"n": A.n, "nnz": A.nnz, "F": int(X.shape[1]), "classes": int(C)}
        os.makedirs(os.path.join(d, "csvs"))
        cmd = [REF_EXE, "-P", str(P), "-R", "1", "-E", str(epochs), "train", ds + "/", str(len(hidden))] + [str(h) for h in hidden]
        t1 = time.time()
        try:
            r = subprocess.run(cmd, cwd=d, capture_output=True, text=True, timeout=3000)
        except subprocess.TimeoutExpired:
            return None, {"why": "reference binary timed out"}
        wall = time.time() - t1
        ep = []
        for line in r.stderr.splitlines():
            f = line.split()
            if len(f) == 4 and f[0].isdigit():
                try:
                    ep.append((int(f[0]), float(f[1]), float(f[2]), float(f[3])))
                except ValueError:
                    pass
        if r.returncode != 0 or len(ep) != epochs:
            return None, {"why": "reference binary failed: " + (r.stderr or r.stdout)[-400:].replace("\n", " | ")}
        info = {"cmd": " ".join(["mg_gcn_ref"] + cmd[1:6] + ["<dataset>"] + cmd[7:]), "wall_s": wall, "dataset_s": t1 - t0,
                "N": meta["n"], "nnz": meta["nnz"], "F": meta["F"], "classes": meta["classes"],
                "dataset_writer": "mggcn_gen (stand-alone executable, no libmggcn.so)" if standalone else "in-process (mg.write_dataset)"}
        try:
            json.dump({"losses": [l for (_, l, _, _) in ep], "acc": [a for (_, _, a, _) in ep], "cmd": info["cmd"]},
                      open(curve_cache_path(shape, scale, P, hidden), "w"))
        except OSError:
            pass
        return ep, info


def reference_curve(mg, shape, scale, P, hidden, epochs):
    """The reference's first `epochs` losses at the same P: from the cache the reference arm leaves behind on this
    box, else from a run of the reference binary (checker use), else -- small graphs only -- from the oracle port."""
    path = curve_cache_path(shape, scale, P, hidden)
    try:
        c = json.load(open(path))
        if len(c["losses"]) >= epochs:
            return c["losses"][:epochs], "oracle/_ref/mg_gcn_ref (cached from the reference arm on this box): " + c.get("cmd", "")
    except (OSError, ValueError, KeyError):
        pass
    ep, info = run_reference_binary(mg, shape, scale, P, hidden, epochs)
    if ep is not None:
        return [l for (_, l, _, _) in ep], "oracle/_ref/mg_gcn_ref: " + info["cmd"]
    why = info["why"]
    n = mg.padded_n(shape, scale)
    if n <= 300000 and mg.shape_params(shape, scale)[1] <= 40e6:
        from oracle import oracle as orc
        A, X, Y, C = mg.generate(shape, seed=SEED, scale=scale, permute=shape not in mg.UNPERMUTED)
        sizes = [X.shape[1]] + hidden + [mg.pad_classes(C, P)]
        A.normalize(True)
        A_T = A.transpose()
        ref = orc.Gcn(sizes, A.n, (A_T.rowptr, A_T.colidx, A_T.vals), (A.rowptr, A.colidx, A.vals), P=P, cuts=mg.row_cuts(A.n, P))
        out = []
        for _ in range(epochs):
            l, _a = ref.train_forward(X, Y)
            ref.backward()
            ref.adam_update(**ADAM)
            out.append(float(l))
        return out, f"oracle port (orc_gcn, P={P}) -- the reference binary was not usable: {why}"
    return None, why


def run_reference(args, mg, rank):
    """The reference arm: the UNMODIFIED reference program (oracle/_ref/mg_gcn_ref, built from
    /root/reference/src where it lies) through its own CLI on the same generated dataset.  Its only
    implementation is CUDA (cuSPARSE/cuBLAS/NCCL); there is no CPU path in the reference, so when the
    binary or a GPU is missing the arm falls back to the oracle port on the host cores."""
    if rank != 0:
        return
    P = args.gpus
    hidden = hidden_of(args, mg)
    E = args.warmup + args.steps
    base = {"metric": "gcn_epoch_ms", "unit": "ms/epoch", "n_gpus": P, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "impl": "reference"}
    dataset = None
    n_raw, edges, f_raw, classes, *_ = mg.shape_params(args.shape, args.scale)
    small_enough = not (args.shape in REF_UNSUPPORTED and args.scale == 1.0)
    ep, info = None, {"why": REF_UNSUPPORTED.get(args.shape, "")}
    if small_enough and os.path.exists(GEN_EXE) and os.path.exists(REF_EXE) and have_gpus(P):
        # the clean path: dataset from the stand-alone writer, reference binary as a subprocess -- libmggcn.so is never
        # loaded into this process
        ep, info = run_reference_binary(mg, args.shape, args.scale, P, hidden, E, None)
        if ep is not None:
            N, nnz, F, C = info["N"], info["nnz"], info["F"], info["classes"]
    if ep is None and small_enough:
        dataset = mg.generate(args.shape, seed=SEED, scale=args.scale, permute=args.shape not in mg.UNPERMUTED)
        A, X, Y, C = dataset
        N, nnz, F = A.n, A.nnz, X.shape[1]
        why_first = info.get("why")
        ep, info = run_reference_binary(mg, args.shape, args.scale, P, hidden, E, dataset)
        if ep is None and why_first and "why" in info and not info["why"]:
            info["why"] = why_first
    elif ep is None:
        N, nnz, F, C = mg.padded_n(args.shape, args.scale), None, (f_raw + 7) // 8 * 8, classes
    base["config"] = {"workload": workload_text(args.shape, N, nnz, F, hidden, C, args.scale), "P": P}
    if ep is not None:
        times = [t for (_, _, _, t) in ep[args.warmup:]]
        ms = float(np.mean(times)) * 1e3
        base.update({"value": ms, "ms_per_step": ms,
                     "e2e": {"value": ms, "unit": "ms/epoch", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                     "cpu_baseline": {"value": ms, "unit": "ms/epoch", "cores": 1, "kind": "reference",
                                      "sample": "the reference's own CUDA build (cuSPARSE/cuBLAS/NCCL; it has no CPU path), "
                                                f"all {E} epochs, mean of the last {args.steps}; one host thread drives {P} GPU(s)"},
                     "reference_detail": {"cmd": info["cmd"], "dataset_writer": info.get("dataset_writer"),
                                          "libmggcn_mapped_in_this_process": maps_libmggcn(), "median_ms": float(np.median(times)) * 1e3,
                                          "min_ms": float(np.min(times)) * 1e3, "losses": [l for (_, l, _, _) in ep][:12],
                                          "wall_s": info["wall_s"]}})
        emit(base)
        return
    base["reference_binary_unusable"] = info["why"]
    if not small_enough:
        # nothing of the reference runs at this size, and the whole graph does not exist on one host either: the
        # bounded CPU sample is the oracle's SpMM on a row slab of the graph against a full-height B
        slab_rows = max(8, int(N * 0.001))
        slab, _, _, _ = mg.generate_rows(args.shape, 0, slab_rows, seed=SEED, permute=False, scale=args.scale, features=False)
        deg = np.full(N, 30, np.uint32)
        deg[:slab_rows] = np.diff(slab.rowptr.astype(np.int64)).astype(np.uint32)
        fwd, bwd = mg.slab_operands(slab, deg, 0)
        sizes = [F] + hidden + [mg.pad_classes(C, P)]
        try:
            cb = cpu_epoch_sample(fwd, bwd, sizes, 1.0, 1, dense_rows=8 << 20)
            cb["value"] *= N / slab_rows
            cb["sample"] = (f"row slab [0, {slab_rows}) of {N} (x{N / slab_rows:.0f}), B wrapped to {8 << 20} rows: " + cb["sample"])
        except MemoryError:
            cb = {"value": None, "unit": "ms/epoch", "cores": 0, "kind": "port", "sample": "skipped: host RAM (B is 57 GB)"}
    else:
        A.normalize(True)
        A_T = A.transpose()
        sizes = [F] + hidden + [mg.pad_classes(C, P)]
        cb = cpu_epoch_sample(A_T, A, sizes, args.cpu_sample, args.cpu_repeats, args.cpu_seconds)
    base.update({"value": cb["value"], "ms_per_step": cb["value"], "cpu_baseline": cb,
                 "e2e": {"value": cb["value"], "unit": "ms/epoch", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    emit(base)


# ------------------------------------------------------------------------------------------------
DEFAULT_TUNE = {}   # library defaults of the knobs --e2e-sweep restores, where they are not 0


def pin_to_gpu_numa_node(local):
    """Pinned staging buffers should live on the NUMA node the GPU hangs off (first touch by this process):
    restrict the process to the CPUs NVML reports as local to the GPU.  Best effort."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


def sampled_layer0_check(mg, eng, args, slab, fwd, sizes, r0, n_rows=48):
    """Full-size parity without a full-size reference: layer 0's output on sampled local rows, recomputed by the
    oracle from first principles -- the feature rows of every neighbour (generated from the counter-based stream,
    wherever the neighbour lives), the reference's initial weights, fp64 accumulation.  Covers the exchange (remote
    neighbours), the SpMM, the GEMM and their fused epilogues at the size that is benchmarked.  Must run before any
    parameter update.  Returns max |got - want| / max(1, |want|_inf per row)."""
    from oracle import oracle as orc
    rng = np.random.default_rng(7)
    n = slab.n
    rows = np.unique(rng.choice(n, size=min(n_rows, n), replace=False))
    deg = np.diff(fwd.rowptr.astype(np.int64))
    rows = np.unique(np.concatenate([rows, [int(deg.argmax())]]))
    srp = np.zeros(len(rows) + 1, np.uint64)
    srp[1:] = np.cumsum(deg[rows])
    sci = np.concatenate([fwd.colidx[int(fwd.rowptr[r]):int(fwd.rowptr[r + 1])] for r in rows])
    sva = np.concatenate([fwd.vals[int(fwd.rowptr[r]):int(fwd.rowptr[r + 1])] for r in rows])
    ids, inv = np.unique(sci, return_inverse=True)
    Xn = mg.generate_feature_rows(args.shape, ids, seed=SEED, permute=args.shape not in mg.UNPERMUTED, scale=args.scale)
    W0 = eng.get("W0").reshape(sizes[0], sizes[1]).astype(np.float64)
    b0 = eng.get("b0").astype(np.float64)
    act = len(sizes) > 2
    if sizes[1] <= sizes[0]:   # GEMM first: HW = X W + b, Z = act(A HW)    (src/gcn.hpp:556-571)
        HW = (Xn.astype(np.float64) @ W0 + b0).astype(np.float32)
        Z = orc.spmm(srp, inv.astype(np.uint32), sva, HW, acc64=True).astype(np.float64)
    else:                      # SpMM first: AH = A X, Z = act(AH W + b)
        AH = orc.spmm(srp, inv.astype(np.uint32), sva, Xn, acc64=True)
        Z = AH.astype(np.float64) @ W0 + b0
    if act:
        Z = np.where(Z > 0.01 * Z, Z, 0.01 * Z)
    worst = 0.0
    for k, r in enumerate(rows):
        got = eng.get_rows("H0", int(r), int(r) + 1)[0].astype(np.float64)
        worst = max(worst, float(np.abs(got - Z[k]).max() / max(1.0, np.abs(Z[k]).max())))
    return {"rows": int(len(rows)), "max_rel_err": worst, "tolerance": 1e-5, "ok": bool(worst <= 1e-5),
            "what": "layer-0 output H0 on sampled rows of rank 0 vs the oracle (fp64) from generated neighbour features"}


def main():
    args = parse()
    claim_stdout()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    numa_cpus = None
    if world > 1:
        numa_cpus = pin_to_gpu_numa_node(local)
        # every rank pre-processes its slab: share the host cores instead of oversubscribing them (torchrun's default
        # of one OpenMP thread makes the host prep 8x slower)
        if os.environ.get("OMP_NUM_THREADS", "1") == "1":
            avail = len(os.sched_getaffinity(0))
            share = max(1, (os.cpu_count() or 8) // world)
            os.environ["OMP_NUM_THREADS"] = str(max(1, min(avail, share) if numa_cpus else share))
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
    ctl = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        ctl = dist.new_group(backend="gloo")  # host-side waits (a NCCL barrier would spin on the GPU meanwhile)
    P = world
    hidden = hidden_of(args, mg)
    permute = args.shape not in mg.UNPERMUTED

    def allsum(vals):
        t = torch.tensor(vals, dtype=torch.float64, device="cuda")
        if P > 1:
            dist.all_reduce(t)
        return [float(x) for x in t]

    # ---- dataset and host pre-processing (src/main.cpp:82-85,139-149), sharded: every rank builds only its slab
    t0 = time.time()
    N = mg.padded_n(args.shape, args.scale)
    cuts = mg.row_cuts(N, P)
    r0, r1 = int(cuts[rank]), int(cuts[rank + 1])
    slab, Xl_np, Yl_np, C = mg.generate_rows(args.shape, r0, r1, seed=SEED, permute=permute, scale=args.scale)
    t_gen = time.time() - t0
    nmax = int(max(cuts[i + 1] - cuts[i] for i in range(P)))
    dl = torch.zeros(nmax, dtype=torch.int32, device="cuda")
    dl[:r1 - r0] = torch.from_numpy(np.diff(slab.rowptr.astype(np.int64)).astype(np.int32)).cuda()
    if P > 1:
        dall = torch.empty(nmax * P, dtype=torch.int32, device="cuda")
        dist.all_gather_into_tensor(dall, dl)
        dall = dall.cpu().numpy().reshape(P, nmax)
        deg_all = np.concatenate([dall[i, :int(cuts[i + 1] - cuts[i])] for i in range(P)]).astype(np.uint32)
    else:
        deg_all = dl.cpu().numpy().astype(np.uint32)
    nnz = int(allsum([slab.nnz])[0])
    F_in = Xl_np.shape[1]
    sizes = [F_in] + hidden + [mg.pad_classes(C, P)]
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
    eng, fwd, bwd = mg.build_rank_engine_sharded(slab, deg_all, sizes, world=P, rank=rank, device=local, cuts=cuts,
                                                 use_graph=not args.no_graph, nccl_id=nccl_id, comm=comm,
                                                 overlap=not args.no_overlap)
    if P > 1 and comm != mg.COMM_NCCL:
        h = eng.comm_handle()
        ht = torch.frombuffer(bytearray(h), dtype=torch.uint8).cuda()
        allh = [torch.empty_like(ht) for _ in range(P)]
        dist.all_gather(allh, ht)
        eng.comm_connect([bytes(t.cpu().numpy().tobytes()) for t in allh])
    Xl = torch.from_numpy(Xl_np).pin_memory()
    Yl = torch.from_numpy(Yl_np).pin_memory()
    eng.set_inputs(Xl, Yl)
    prep_s = time.time() - t0
    log(f"rank {rank}: slab {r1 - r0} rows, {slab.nnz} entries, generated in {t_gen:.1f} s, prep {prep_s:.1f} s")
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

    # ---- parity: sampled rows of layer 0 at full size, then 10 epochs from the reference initialisation
    parity = None
    if not args.skip_parity:
        parity = {}
        eng.train_forward()
        if rank == 0:
            try:
                parity["sampled_rows"] = sampled_layer0_check(mg, eng, args, slab, fwd, sizes, r0)
            except Exception as ex:  # the checker must never cost the line
                parity["sampled_rows"] = {"error": repr(ex)}
        curve = []
        for _ in range(PARITY_EPOCHS):
            eng.epoch(**ADAM)
            ls, cs = eng.loss()
            tot = allsum([ls, cs])
            curve.append(tot[0] / N)
        parity["loss_curve"] = curve
        barrier()
        if rank == 0:
            ref, src = reference_curve(mg, args.shape, args.scale, P, hidden, PARITY_EPOCHS)
            parity["reference_curve"] = ref
            parity["reference_source"] = src
            if ref is not None:
                d = float(np.max(np.abs(np.array(curve) - np.array(ref))))
                parity.update({"max_abs_diff": d, "tolerance": 1e-4, "ok": bool(d <= 1e-4)})
            else:
                sr = parity.get("sampled_rows", {})
                parity.update({"max_abs_diff": None, "tolerance": 1e-4, "ok": bool(sr.get("ok", False)) and all(np.isfinite(curve)),
                               "note": "no reference curve at this size; ok = sampled-row check green and finite losses"})
            log("parity:", json.dumps(parity))
        if P > 1:
            dist.barrier(group=ctl)

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
    # the reference's loop reads the loss back and synchronises every epoch (src/main.cpp:158-169): same here
    ms_sync_dev, ms_sync_wall = timed(lambda: (eng.epoch(**ADAM), eng.loss()), args.steps)

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
    # the copy alone, all ranks at once: the PCIe rate an epoch has to hide
    barrier()
    h0 = time.perf_counter()
    for _ in range(3):
        eng.set_inputs(Xl, Yl, async_=1)
    eng.sync()
    h2d_s = (time.perf_counter() - h0) / 3
    h2d_bytes_rank = Xl.numel() * 4 + Yl.numel() * 4
    # what the input copy costs when every rank copies at once: slowest and fastest rank (the ranks share PCIe
    # switches and one host memory system), and the same pipelined step with the copy done by a few CTAs reading
    # the pinned rows over PCIe (mggcn_tune "h2d_sm") instead of by a copy engine
    e2e_diag = None
    if P > 1 or args.e2e_variants:
        t = torch.tensor([h2d_s, -h2d_s], device="cuda", dtype=torch.float64)
        if P > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_diag = {"h2d_alone_GBps_slowest_rank": h2d_bytes_rank / float(t[0]) / 1e9,
                    "h2d_alone_GBps_fastest_rank": h2d_bytes_rank / -float(t[1]) / 1e9,
                    "step_ms": {"default (pieces by size, see h2d_piece_bytes)": e2e_wall}}
        tuned = dict(kv.split("=") for kv in args.tune)
        for key, val in (("h2d_chunk_kb", 1 << 22), ("h2d_chunk_kb", 128), ("h2d_chunk_kb", 512)):
            mg.lib().mggcn_tune(key.encode(), val)
            eng.set_inputs(Xl, Yl, async_=2)
            e2e_step()
            _, w = timed(e2e_step, args.steps)
            e2e_diag["step_ms"][f"{key}={val}"] = w
            mg.lib().mggcn_tune(key.encode(), int(tuned.get(key, 0)))
        # the two sides of the pipelined step, loaded by each other: a foreign stream keeps copying X from pinned
        # memory while the resident-input epoch is timed, and the copies are timed while the epochs run
        try:
            cs = torch.cuda.Stream()
            dst = torch.empty(Xl.shape, dtype=Xl.dtype, device="cuda")
            ncp = 2 * args.steps
            evs = [torch.cuda.Event(enable_timing=True) for _ in range(ncp + 1)]
            barrier()
            with torch.cuda.stream(cs):
                evs[0].record(cs)
                for i in range(ncp):
                    dst.copy_(Xl, non_blocking=True)
                    evs[i + 1].record(cs)
            ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ea.record(sc)
            for _ in range(args.steps):
                eng.epoch(**ADAM)
            eb.record(sc)
            eng.sync()
            torch.cuda.synchronize()
            t = torch.tensor([ea.elapsed_time(eb) / args.steps,
                              float(np.median([evs[i].elapsed_time(evs[i + 1]) for i in range(max(1, args.steps // 2))]))],
                             device="cuda", dtype=torch.float64)
            if P > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_diag["loaded"] = {"epoch_ms_while_copying": float(t[0]), "copy_ms_while_computing_slowest_rank": float(t[1]),
                                  "copy_ms_alone_slowest_rank": e2e_diag and h2d_bytes_rank / (e2e_diag["h2d_alone_GBps_slowest_rank"] * 1e9) * 1e3}
            del dst
        except Exception as ex:  # diagnostics only
            e2e_diag["loaded"] = {"error": repr(ex)}
        # combinations of exchange / copy knobs, each timed with resident inputs and end to end (the captured epoch
        # is re-recorded when an exchange knob changes)
        for combo in args.e2e_sweep:
            kv = [x.split("=") for x in combo.split(",")]
            for k, v in kv:
                mg.lib().mggcn_tune(k.encode(), int(v))
            for _ in range(3):
                eng.epoch(**ADAM)
            dv, _ = timed(lambda: eng.epoch(**ADAM), args.steps)
            eng.set_inputs(Xl, Yl, async_=2)
            e2e_step()
            _, w = timed(e2e_step, args.steps)
            e2e_diag.setdefault("sweep", {})[combo] = {"value_ms": dv, "e2e_ms": w}
            for k, v in kv:
                mg.lib().mggcn_tune(k.encode(), int(tuned.get(k, DEFAULT_TUNE.get(k, 0))))
        for _ in range(3):
            eng.epoch(**ADAM)
        eng.set_inputs(Xl, Yl, async_=2)
        e2e_step()

    # ---- dominant kernel, live: eager pass with the engine's named timers (CUDA events on its stream)
    eng.set_mode(use_graph=False, timers=True)
    per = {}
    for _ in range(3):
        eng.epoch(**ADAM)
        eng.sync()
        for line in eng.timers().splitlines():
            name, ms = line.rsplit(":", 1)
            per.setdefault(name, []).append(float(ms))
    cfgs = eng.spmm_configs()
    eng.set_mode(use_graph=not args.no_graph, timers=False)
    med = {k: float(np.median(v)) for k, v in per.items()}
    # the dominant launches: forward aggregations at the widest aggregated width
    n_loc = r1 - r0
    L = len(sizes) - 1
    widths = [min(sizes[l], sizes[l + 1]) for l in range(L)]
    F = max(widths)
    dom_layers = [l for l in range(L) if widths[l] == F]
    fused = P > 1 and args.comm == "p2p"
    if P == 1 or fused:
        keys = [f"{l}_0_matmul-spmm" for l in dom_layers]
        dom_nnz = fwd.nnz
        alg = spmm_alg_bytes(n_loc, N, dom_nnz, F)
        dom_cols = N
        scope = "whole matrix" if P == 1 else f"block row {n_loc}x{N}, own block + one gathered launch"
    else:
        # per-stage launches: the own-block launch (stage / group 0), one n_loc x n_loc block
        keys = [f"{l}_0_0_matmul-spmm-stage" for l in dom_layers]
        blk = fwd.block_rows(np.array([0, n_loc], np.uint32), cuts, only_row=0)[rank]
        dom_nnz = blk.nnz
        alg = spmm_alg_bytes(n_loc, n_loc, dom_nnz, F)
        dom_cols = n_loc
        scope = f"own {n_loc}x{n_loc} block (first launch of the aggregation)"
    dom = [med[k] for k in keys if k in med]
    dom_ms = float(np.mean(dom)) if dom else float("nan")
    dom_cfg = next((cfgs[k] for k in keys if k in cfgs), "unknown")
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "spmm_traffic.json"))).get(f"{args.shape}_P{P}_F{F}")
    except (OSError, ValueError):
        pass
    achieved = alg / (dom_ms * 1e-3) / 1e9 if dom_ms == dom_ms and dom_ms > 0 else None
    # second roofline: rows of B are gathered avg-degree times, and when B fits the L2 that stream (nnz*F*4 bytes)
    # is bounded by the L2 -> SM read bandwidth (measured on B200 by tools/l2_peak.py, kept in profiles/l2_peak.json)
    l2_roof = None
    try:
        l2pk = json.load(open(os.path.join(ROOT, "profiles", "l2_peak.json")))
        if dom_ms == dom_ms and dom_ms > 0 and l2pk.get("read_GBps"):
            gat = dom_nnz * F * 4 / (dom_ms * 1e-3) / 1e9
            l2_roof = {"bound": "l2", "achieved": gat, "peak": float(l2pk["read_GBps"]), "unit": "GB/s",
                       "frac": gat / float(l2pk["read_GBps"]), "peak_source": "profiles/l2_peak.json (tools/l2_peak.py on B200)",
                       "applies": bool(dom_cols * F * 4 <= 126e6),
                       "note": "gather stream nnz*F*4 over the launch time against the measured L2->SM read bandwidth; "
                               "applies when B (cols*F*4 bytes) is L2-resident, else the gathers are DRAM reads"}
    except (OSError, ValueError):
        pass
    spmm_ms = sum(v for k, v in med.items() if k.endswith("matmul-spmm"))
    gemm_ms = sum(v for k, v in med.items() if k.endswith("matmul-gemm"))
    xfer_ms = sum(v for k, v in med.items() if k.endswith("matmul-bcast"))

    # ---- exposed communication (north_star: < 10 % of the epoch): the same epoch with the peer copies left out
    # (barriers, events and launches stay; results are garbage, so this runs last)
    exposed = None
    if P > 1 and args.comm != "nccl" and args.comm != "p2p" and not args.skip_exposed:
        mg.lib().mggcn_tune(b"skip_exchange", 1)
        for _ in range(3):
            eng.epoch(**ADAM)
        ms_noex, _ = timed(lambda: eng.epoch(**ADAM), args.steps)
        mg.lib().mggcn_tune(b"skip_exchange", 0)
        for _ in range(2):
            eng.epoch(**ADAM)
        ms_again, _ = timed(lambda: eng.epoch(**ADAM), args.steps)
        exposed = {"epoch_ms_with_exchange": ms_again, "epoch_ms_without_copies": ms_noex,
                   "exposed_ms": ms_again - ms_noex, "exposed_frac": (ms_again - ms_noex) / ms_again if ms_again > 0 else None,
                   "how": "mggcn_tune('skip_exchange', 1): peer copies not enqueued, everything else identical; both "
                          "timed back to back after the main measurement"}

    if rank != 0:
        eng.close()
        if P > 1:
            dist.barrier(group=ctl)
            dist.destroy_process_group()
        return

    cpu = None
    if P == 1 and not args.skip_cpu:
        cpu = cpu_epoch_sample(fwd, bwd, sizes, args.cpu_sample, args.cpu_repeats, args.cpu_seconds)

    out = {
        "metric": "gcn_epoch_ms", "value": ms_dev, "unit": "ms/epoch", "n_gpus": P, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_dev, "higher_is_better": False, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_text(args.shape, N, nnz, F_in, hidden, C, args.scale), "P": P},
        "run": {"partition": "1-D row blocks, replicated weights", "comm": args.comm if P > 1 else "none",
                "graph": not args.no_graph, "overlap": not args.no_overlap, "scale": args.scale, "sizes": sizes,
                "classes_padded_to": sizes[-1], "l2": "inputs larger than L2: no flush needed",
                **({"tune": args.tune} if args.tune else {})},
        "parity": parity,
        "e2e": {"value": e2e_wall, "unit": "ms/epoch", "device_ms": e2e_dev,
                "h2d_bytes_per_step": int(N * F_in * 4 + N * 4), "d2h_bytes_per_step": 16 * P,
                "h2d_GBps_per_rank_all_ranks_copying": h2d_bytes_rank / h2d_s / 1e9, "numa_local_cpus": numa_cpus,
                **({"diag": e2e_diag} if e2e_diag else {})},
        "synced_epoch": {"value": ms_sync_wall, "device_ms": ms_sync_dev,
                         "note": "epoch + loss read-back + host sync every step, like the reference's loop (src/main.cpp:158-169)"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": dom_cfg, "scope": f"forward aggregation, F={F}, {scope}", "achieved": achieved,
                     "peak": peak,
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s (B200_PROFILING.md)",
                     "unit": "GB/s", "frac": (achieved / peak) if achieved else None, "traffic": traffic,
                     "alg_bytes_per_launch": alg, "launch_ms": dom_ms,
                     "gather_GBps": (dom_nnz * F * 4 / (dom_ms * 1e-3) / 1e9) if dom_ms > 0 else None,
                     "GFLOPs": (2 * dom_nnz * F / (dom_ms * 1e-3) / 1e9) if dom_ms > 0 else None,
                     "l2": l2_roof,
                     "note": "achieved/frac use the algorithmic bytes of ONE pass over the operands (SURVEY.md 8d) and "
                             "launch_ms is the whole aggregation (all its windows / blocks); gather_GBps = nnz*F*4/t is "
                             "the operand traffic the L2 has to serve when B is gathered avg-degree times"},
        "breakdown_ms": {"spmm": spmm_ms, "gemm": gemm_ms, "exchange": xfer_ms,
                         "loss": sum(v for k, v in med.items() if k.endswith("loss-layer")),
                         "adam": med.get("adam-update"), "allreduce": med.get("allreduce"),
                         "eager_epoch_sum": sum(v for k, v in med.items() if not k.endswith(("-stage", "-bcast")))},
        "wall_ms_per_step": ms_wall, "prep_s": prep_s, "gen_s": t_gen,
        "loss": parity["loss_curve"][-1] if parity else None,
    }
    if P > 1:
        # exchange against the NVLink roofline (BASELINE.json north_star): bytes this rank receives per epoch =
        # sum over the aggregations of (P-1)/P * N * ld * 4, over the time its copies are in flight (eager timers)
        try:
            ws = [a for k, a in layer_plan(sizes) if k == "spmm"]
            rx = sum((P - 1) / P * N * ((w + 3) // 4 * 4) * 4 for w in ws)
            out["exchange"] = {"rx_bytes_per_rank_per_epoch": int(rx), "copy_ms_per_epoch": xfer_ms,
                               "GBps": (rx / (xfer_ms * 1e-3) / 1e9) if xfer_ms > 0 else None,
                               "nvlink_peak_GBps": 900.0,
                               "nvlink_frac": (rx / (xfer_ms * 1e-3) / 1e9 / 900.0) if xfer_ms > 0 else None,
                               **(exposed or {})}
        except Exception as ex:  # accounting only: never lose the line
            out["exchange"] = {"error": repr(ex)}
    if cpu:
        out["cpu_baseline"] = cpu
    emit(out)
    eng.close()
    if P > 1:
        dist.barrier(group=ctl)
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
