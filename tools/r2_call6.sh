# 1-GPU call: whole GPU suite on the build with the pre-split weight operand, bench lines (reddit, products),
# GEMM variants, launch list + one --set full capture of the shipped SpMM launches
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -rs --timeout=900 -p no:cacheprovider > gpurun_out/r2f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2f_pytest.log
tail -n 4 gpurun_out/r2f_pytest.log
python tools/gemm_bench.py --presplit 0 1 --split 0 1 > gpurun_out/r2f_gemm_bench.jsonl 2>&1; echo "gemm_bench rc=$?"
python bench.py --impl reference > gpurun_out/r2f_ref_reddit_n1.json 2> gpurun_out/r2f_ref_reddit_n1.err
python bench.py > gpurun_out/r2f_bench_reddit_n1.json 2> gpurun_out/r2f_bench_reddit_n1.err; echo "bench reddit rc=$?"
python bench.py --skip-parity --skip-cpu --tune gemm_presplit=0 > gpurun_out/r2f_bench_reddit_n1_nopresplit.json 2> gpurun_out/r2f_bench_reddit_n1_nopresplit.err
python bench.py --skip-parity --skip-cpu --tune gemm_split=1 > gpurun_out/r2f_bench_reddit_n1_trunc.json 2> gpurun_out/r2f_bench_reddit_n1_trunc.err
python - <<'PY'
import json
for t in ("", "_nopresplit", "_trunc"):
    try:
        d = json.load(open(f"gpurun_out/r2f_bench_reddit_n1{t}.json"))
        print(t or "default", "value %.3f e2e %.3f gemm %.3f spmm %.3f parity %s clocks %s" % (d["value"], d["e2e"]["value"], d["breakdown_ms"]["gemm"], d["breakdown_ms"]["spmm"], (d.get("parity") or {}).get("ok"), d["clocks"]["sm_mhz"]))
    except Exception as ex:
        print(t, "no line:", ex)
PY
python bench.py --shape products --impl reference > gpurun_out/r2f_ref_products_n1.json 2> gpurun_out/r2f_ref_products_n1.err
python bench.py --shape products > gpurun_out/r2f_bench_products_n1.json 2> gpurun_out/r2f_bench_products_n1.err; echo "bench products rc=$?"
head -c 300 gpurun_out/r2f_bench_products_n1.json; echo
CMD="python bench.py --steps 2 --warmup 1 --skip-parity --skip-cpu"
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2f_launches_n1.csv $CMD > gpurun_out/r2f_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
CMD2="python tools/spmm_bench.py --shape reddit --F 128 --variants auto:"
$CMD2 > gpurun_out/r2f_plain_spmm.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:spmm_kernel -s 4 -c 2 -o gpurun_out/r2f_prof_spmm_reddit $CMD2 > gpurun_out/r2f_ncu_spmm.log 2>&1; echo "ncu spmm rc=$?"
ls -la gpurun_out | tail -n 8
