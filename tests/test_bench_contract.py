"""bench.py's output contract on a machine without a GPU: the reference arm falls back to the oracle port on the
host cores (the reference program itself is CUDA-only), prints exactly ONE JSON line on stdout and carries the keys
the driver reads.  Our own arm must refuse to run without a GPU instead of falling back."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(*args):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                          timeout=600, cwd=ROOT)


def have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_reference_arm_prints_one_json_line():
    r = run("--impl", "reference", "--scale", "0.01", "--steps", "1", "--warmup", "1", "--cpu-repeats", "1")
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "gcn_epoch_ms" and d["unit"] == "ms/epoch"
    assert d["higher_is_better"] is False and d["scaling"] == "strong" and d["vs_baseline"] is None
    assert d["value"] > 0 and d["ms_per_step"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "ms/epoch", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["sample"] and cb["value"] == d["value"]
    assert "workload" in d["config"] and "model" not in d["config"]


def test_accounting_helpers():
    sys.path.insert(0, ROOT)
    import bench
    # every operand once
    assert bench.spmm_alg_bytes(10, 20, 100, 8) == 100 * 8 + 11 * 4 + 20 * 8 * 4 + 10 * 8 * 4
    assert bench.spmm_alg_bytes(10, 20, 100, 8, beta_nonzero=True, w=8) == 100 * 8 + 11 * 8 + 20 * 8 * 4 + 2 * 10 * 8 * 4
    # reddit 608 -> 128 -> 128 -> 128 -> 41: 4 forward + 3 backward aggregations (layer 0's is skipped), widths as in
    # SURVEY.md 8(a)
    ops = bench.layer_plan([608, 128, 128, 128, 41])
    assert [a for k, a in ops if k == "spmm"] == [128, 128, 128, 41, 41, 128, 128]
    # products 100 -> 256 -> 256 -> 256 -> 47: the first layer aggregates before it transforms
    ops = bench.layer_plan([100, 256, 256, 256, 47])
    assert [a for k, a in ops if k == "spmm"][:4] == [100, 256, 256, 47]


@pytest.mark.skipif(have_gpu(), reason="checks the no-GPU refusal")
def test_own_arm_refuses_to_run_without_a_gpu():
    r = run("--scale", "0.01", "--steps", "1", "--warmup", "1", "--skip-cpu")
    assert r.returncode != 0
    assert "needs a GPU" in r.stderr
    assert not [l for l in r.stdout.splitlines() if l.strip().startswith("{")]


def test_standalone_dataset_writer_matches_the_library(mg, tmp_path):
    """bench.py's reference arm writes its dataset with mg-gcn_b200/cpp/bin/mggcn_gen (no libmggcn.so in that
    process); the files must be the ones the repo arm's in-process generator writes, byte for byte."""
    import subprocess
    exe = os.path.join(ROOT, "mg-gcn_b200", "cpp", "bin", "mggcn_gen")
    if not os.path.exists(exe):
        r = subprocess.run(["make", "-C", os.path.join(ROOT, "mg-gcn_b200", "cpp"), "-j4"], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[-2000:]
    ldd = subprocess.run(["ldd", exe], capture_output=True, text=True).stdout
    assert "libmggcn" not in ldd and "libcudart" not in ldd
    for shape, permute in (((3000, 40000, 50, 7, 0, 0.0, 0), 1), ((2500, 60000, 33, 41, 1, 2.2, 300), 0)):
        n_raw, edges, f_raw, classes, model, gamma, max_deg = shape
        a, b = str(tmp_path / f"exe{permute}"), str(tmp_path / f"lib{permute}")
        r = subprocess.run([exe, str(n_raw), str(edges), str(f_raw), str(classes), str(model), repr(gamma), str(max_deg), "1",
                            str(permute), a], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        meta = json.loads(r.stdout.strip().splitlines()[-1])
        A, X, Y, C = mg.generate(shape, seed=1, permute=bool(permute))
        mg.write_dataset(b, A, X, Y)
        assert meta == {"n": A.n, "nnz": A.nnz, "F": X.shape[1], "classes": C}
        for f in ("graph.bin", "features.bin", "labels.bin", "sets.bin"):
            assert open(os.path.join(a, f), "rb").read() == open(os.path.join(b, f), "rb").read(), f
