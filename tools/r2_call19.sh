# 8-GPU call: reddit-shape line with the new defaults (input transfer in pieces, one stream per small block) and the
# end-to-end variants beside it; 2-GPU surface test on an 8-GPU box
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port"
t0=$(date +%s)
timeout 120 $TR 29831 bench.py --gpus 8 --skip-cpu --skip-parity > gpurun_out/r2s_bench_reddit_n8.json 2> gpurun_out/r2s_bench_reddit_n8.err; echo "reddit rc=$? t=$(( $(date +%s) - t0 ))"
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/r2s_bench_reddit_n8.json"))
    x = d["exchange"]
    print("value %.3f synced %.3f e2e %.3f exposed %.3f (%.3f) copy GB/s %.0f" % (d["value"], d["synced_epoch"]["value"], d["e2e"]["value"], x["exposed_ms"], x["exposed_frac"], x["GBps"]))
    print(json.dumps(d["e2e"].get("diag"), indent=1))
    print(d["breakdown_ms"])
except Exception as ex:
    print("no line:", ex)
PY
for k in 1 2; do timeout 60 mg-gcn_b200/cpp/bin/test_surface dist > gpurun_out/r2s_dist_$k.log 2>&1; echo "surface dist run $k rc=$? fails=$(grep -c 'EXPECT failed' gpurun_out/r2s_dist_$k.log)"; grep "layer stack" gpurun_out/r2s_dist_$k.log | head -6; done
