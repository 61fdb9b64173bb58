# 8-GPU call: two world-8 checker runs (uneven cuts; one launch per stage), reddit-shape line with the end-to-end copy
# diagnostics, products-shape with the shape-chosen stage groups, papers100M-shape at full size (BASELINE.json config 5)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port"
t0=$(date +%s)
timeout 150 python -m pytest tests/test_gpu_engine.py -m gpu -q -rs --timeout=140 -p no:cacheprovider -k "eight_gpu and (extra4 or extra5)" > gpurun_out/r2o_pytest_n8.log 2>&1; echo "pytest rc=$? t=$(( $(date +%s) - t0 ))"
tail -n 3 gpurun_out/r2o_pytest_n8.log
timeout 120 $TR 29803 bench.py --gpus 8 --skip-cpu --skip-parity > gpurun_out/r2o_bench_reddit_n8.json 2> gpurun_out/r2o_bench_reddit_n8.err; echo "reddit rc=$? t=$(( $(date +%s) - t0 ))"
timeout 120 $TR 29802 bench.py --gpus 8 --shape products --skip-cpu --skip-parity > gpurun_out/r2o_bench_products_n8.json 2> gpurun_out/r2o_bench_products_n8.err; echo "products rc=$? t=$(( $(date +%s) - t0 ))"
timeout 330 $TR 29801 bench.py --gpus 8 --shape papers100M --steps 3 --skip-cpu > gpurun_out/r2o_bench_papers100M_n8.json 2> gpurun_out/r2o_bench_papers100M_n8.err; echo "papers100M rc=$? t=$(( $(date +%s) - t0 ))"
grep -E "rank 0|parity|Error|error" gpurun_out/r2o_bench_papers100M_n8.err | tail -n 6 | cut -c1-600
python - <<'PY'
import json
for t in ("reddit", "products", "papers100M"):
    try:
        d = json.load(open(f"gpurun_out/r2o_bench_{t}_n8.json"))
        x = d.get("exchange") or {}
        print(t, "value %.3f e2e %.3f" % (d["value"], d["e2e"]["value"]), "diag", d["e2e"].get("diag"), "parity", (d.get("parity") or {}).get("ok"),
              (d.get("parity") or {}).get("max_abs_diff"), "exposed", x.get("exposed_ms"), x.get("exposed_frac"), "bd", d["breakdown_ms"], "prep", d["prep_s"], d["gen_s"])
    except Exception as ex:
        print(t, "no line:", ex)
PY
