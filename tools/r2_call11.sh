# 1-GPU call: GEMM producer issuing one box per operand and k-block, splitter loads hoisted
mkdir -p gpurun_out
python -m pytest tests/test_gpu_dense.py tests/test_gpu_engine.py tests/test_cpp_surface.py -m gpu -q -rs --timeout=900 -p no:cacheprovider > gpurun_out/r2k_pytest.log 2>&1; echo "pytest rc=$?"
tail -n 3 gpurun_out/r2k_pytest.log
for ts in 1 0; do
  timeout 300 python tools/gemm_bench.py --tune gemm_ts=$ts > gpurun_out/r2k_gemm_reddit_ts$ts.jsonl 2>&1; echo "gemm_bench reddit ts=$ts rc=$?"
  timeout 300 python tools/gemm_bench.py --n 2449032 --shapes products --tune gemm_ts=$ts > gpurun_out/r2k_gemm_products_ts$ts.jsonl 2>&1; echo "gemm_bench products ts=$ts rc=$?"
done
timeout 300 python tools/gemm_bench.py --kgroup 4 > gpurun_out/r2k_gemm_reddit_ts1_kg4.jsonl 2>&1
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2k_gemm_*.jsonl")):
    tot = 0; rows = []
    for l in open(f):
        try: d = json.loads(l)
        except Exception: print(f, l[:300]); continue
        if d["tag"].startswith("tcgen05"):
            tot += d["ms"]
            rows.append("%d/%d/%d:%.3f(%.0f,%.1e)" % (d["M"] if d["M"] < 9999 else 0, d["N"], d["K"] if d["K"] < 9999 else 0, d["ms"], d["GBps"], d["max_rel_err"]))
    print(f, "sum %.3f" % tot, " ".join(rows))
PY
python bench.py --skip-cpu --skip-parity > gpurun_out/r2k_bench_reddit_n1.json 2> gpurun_out/r2k_bench_reddit_n1.err
python bench.py --skip-cpu --skip-parity --shape products > gpurun_out/r2k_bench_products_n1.json 2> gpurun_out/r2k_bench_products_n1.err
python - <<'PY'
import json
for t in ("reddit", "products"):
    try:
        d = json.load(open(f"gpurun_out/r2k_bench_{t}_n1.json"))
        print(t, "value %.3f e2e %.3f gemm %.3f spmm %.3f clocks %s" % (d["value"], d["e2e"]["value"], d["breakdown_ms"]["gemm"], d["breakdown_ms"]["spmm"], d["clocks"]["sm_mhz"]))
    except Exception as ex:
        print(t, "no line:", ex)
PY
CMD3="python tools/gemm_bench.py"
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 4 -c 2 -o gpurun_out/r2k_prof_gemm_ts $CMD3 > gpurun_out/r2k_ncu_gemm.log 2>&1; echo "ncu gemm rc=$?"
