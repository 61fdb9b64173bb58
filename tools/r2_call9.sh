# 1-GPU call: GEMM with L2 prefetch ahead of the TMA ring (distance sweep), narrow-lane SpMM at the class widths
mkdir -p gpurun_out
python -m pytest tests/test_gpu_dense.py tests/test_gpu_engine.py -m gpu -q -rs --timeout=900 -p no:cacheprovider > gpurun_out/r2i_pytest.log 2>&1; echo "pytest rc=$?"
tail -n 3 gpurun_out/r2i_pytest.log
for d in 0 4 8 16 32; do
  python tools/gemm_bench.py --tune gemm_prefetch=$d > gpurun_out/r2i_gemm_pf$d.jsonl 2>&1; echo "gemm_bench prefetch=$d rc=$?"
done
python tools/gemm_bench.py --n 2449032 --shapes products --tune gemm_prefetch=0 > gpurun_out/r2i_gemm_products_pf0.jsonl 2>&1
python tools/gemm_bench.py --n 2449032 --shapes products --tune gemm_prefetch=8 > gpurun_out/r2i_gemm_products_pf8.jsonl 2>&1
python tools/gemm_bench.py --n 2449032 --shapes products --tune gemm_prefetch=16 > gpurun_out/r2i_gemm_products_pf16.jsonl 2>&1
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2i_gemm_*.jsonl")):
    tot = 0; rows = []
    for l in open(f):
        try: d = json.loads(l)
        except Exception: continue
        if d["tag"].startswith("tcgen05"): tot += d["ms"]; rows.append("%d/%d/%d:%.3f(%.0f)" % (d["M"] if d["M"] < 9999 else 0, d["N"], d["K"] if d["K"] < 9999 else 0, d["ms"], d["GBps"]))
    print(f, "sum %.3f ms" % tot, " ".join(rows))
PY
python tools/spmm_bench.py --shape reddit --F 44 --variants "auto:" "n1:narrow=1" "n2:narrow=2" > gpurun_out/r2i_spmm_reddit_f44.jsonl 2>&1
python tools/spmm_bench.py --shape products --F 48 --variants "auto:" "n1:narrow=1" "n2:narrow=2" > gpurun_out/r2i_spmm_products_f48.jsonl 2>&1
grep -h ms_med gpurun_out/r2i_spmm_reddit_f44.jsonl gpurun_out/r2i_spmm_products_f48.jsonl | cut -c1-260
python bench.py --skip-parity --skip-cpu > gpurun_out/r2i_bench_reddit_n1.json 2> gpurun_out/r2i_bench_reddit_n1.err
python bench.py --skip-parity --skip-cpu --shape products > gpurun_out/r2i_bench_products_n1.json 2> gpurun_out/r2i_bench_products_n1.err
python - <<'PY'
import json
for t in ("reddit", "products"):
    try:
        d = json.load(open(f"gpurun_out/r2i_bench_{t}_n1.json"))
        print(t, "value %.3f e2e %.3f gemm %.3f spmm %.3f clocks %s" % (d["value"], d["e2e"]["value"], d["breakdown_ms"]["gemm"], d["breakdown_ms"]["spmm"], d["clocks"]["sm_mhz"]))
    except Exception as ex:
        print(t, "no line:", ex)
PY
