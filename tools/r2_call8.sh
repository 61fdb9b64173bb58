# 8-GPU call: papers100M-shape (BASELINE.json config 5), products-shape with the stage grouping chosen by the
# engine against the doubling scheme, world-8 checker on uneven cuts and on both groupings
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29741"
timeout 200 $TR tests/dist_engine_check.py --comm p2p-grouped --skew 11 --xgroups 0 > gpurun_out/r2h_check_skew_auto.log 2>&1; echo "checker skew/auto rc=$?"
timeout 200 $TR tests/dist_engine_check.py --comm p2p-grouped --graph --xgroups 1 > gpurun_out/r2h_check_singles_graph.log 2>&1; echo "checker singles/graph rc=$?"
tail -n 2 gpurun_out/r2h_check_skew_auto.log gpurun_out/r2h_check_singles_graph.log
timeout 600 $TR bench.py --gpus 8 --shape papers100M --steps 5 > gpurun_out/r2h_bench_papers100M_n8.json 2> gpurun_out/r2h_bench_papers100M_n8.err; echo "papers100M rc=$?"
tail -n 3 gpurun_out/r2h_bench_papers100M_n8.err
head -c 2500 gpurun_out/r2h_bench_papers100M_n8.json; echo
timeout 300 $TR bench.py --gpus 8 --shape products --skip-cpu > gpurun_out/r2h_bench_products_n8.json 2> gpurun_out/r2h_bench_products_n8.err; echo "products auto rc=$?"
timeout 200 $TR bench.py --gpus 8 --shape products --skip-cpu --skip-parity --tune xgroups=2 > gpurun_out/r2h_bench_products_n8_doubling.json 2> gpurun_out/r2h_bench_products_n8_doubling.err; echo "products doubling rc=$?"
python - <<'PY'
import json
for t in ("", "_doubling"):
    try:
        d = json.load(open(f"gpurun_out/r2h_bench_products_n8{t}.json")); x = d["exchange"]
        print(t or "auto", "value %.3f e2e %.3f copy_ms %.3f GBps %.0f exposed %.3f parity %s" % (d["value"], d["e2e"]["value"], x["copy_ms_per_epoch"], x["GBps"], x["exposed_ms"], (d.get("parity") or {}).get("ok")))
    except Exception as ex:
        print(t, "no line:", ex)
PY
