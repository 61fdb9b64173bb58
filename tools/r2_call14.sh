# 1-GPU call (re-entry): whole GPU suite + smoke on the shipped build, GEMM micro-bench, reddit / products bench lines,
# ncu launch list of the bench command and a --set full capture of the GEMM
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -rs --timeout=900 -p no:cacheprovider > gpurun_out/r2n_pytest.log 2>&1; echo "pytest rc=$?"
tail -n 12 gpurun_out/r2n_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2n_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/r2n_smoke.log
timeout 300 python tools/gemm_bench.py --kgroup 2 4 8 > gpurun_out/r2n_gemm_reddit.jsonl 2>&1; echo "gemm_bench reddit rc=$?"
timeout 300 python tools/gemm_bench.py --tune gemm_ts=0 > gpurun_out/r2n_gemm_reddit_ts0.jsonl 2>&1; echo "gemm_bench reddit ts0 rc=$?"
timeout 300 python tools/gemm_bench.py --n 2449032 --shapes products --kgroup 2 4 > gpurun_out/r2n_gemm_products.jsonl 2>&1; echo "gemm_bench products rc=$?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2n_gemm_*.jsonl")):
    tot = {}; rows = []
    for l in open(f):
        try: d = json.loads(l)
        except Exception: print(f, l[:300]); continue
        if d["tag"].startswith("tcgen05"):
            tot[d["kgroup"]] = tot.get(d["kgroup"], 0) + d["ms"]
            rows.append("%d/%d/%d kg%d:%.3f(%.0f,%.1e)" % (d["M"] if d["M"] < 9999 else 0, d["N"], d["K"] if d["K"] < 9999 else 0, d["kgroup"], d["ms"], d["GBps"], d["max_rel_err"]))
    print(f.split("/")[-1], "sum", {k: round(v, 3) for k, v in tot.items()}, " ".join(rows))
PY
timeout 600 python bench.py > gpurun_out/r2n_bench_reddit_n1.json 2> gpurun_out/r2n_bench_reddit_n1.err; echo "bench rc=$?"
timeout 600 python bench.py --skip-cpu --shape products > gpurun_out/r2n_bench_products_n1.json 2> gpurun_out/r2n_bench_products_n1.err
python - <<'PY'
import json
for t in ("reddit", "products"):
    try:
        d = json.load(open(f"gpurun_out/r2n_bench_{t}_n1.json"))
        print(t, "value %.3f e2e %.3f gemm %.3f spmm %.3f parity %s %s clocks %s" % (d["value"], d["e2e"]["value"], d["breakdown_ms"]["gemm"], d["breakdown_ms"]["spmm"], d["parity"]["ok"], d["parity"]["max_abs_diff"], d["clocks"]["sm_mhz"]))
    except Exception as ex:
        print(t, "no line:", ex)
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2n_launches.csv python bench.py --steps 2 --warmup 1 --skip-parity --skip-cpu > gpurun_out/r2n_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 4 -c 3 -o gpurun_out/r2n_prof_gemm python tools/gemm_bench.py > gpurun_out/r2n_ncu_gemm.log 2>&1; echo "ncu gemm rc=$?"
