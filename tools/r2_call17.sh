# 8-GPU call: what the pipelined end-to-end step at reddit shape loses (epoch loaded by the input copy, copy loaded by the
# epoch, exchange / copy knob combinations), and the 2-GPU surface test after the replica write-after-read fix
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port"
t0=$(date +%s)
timeout 150 $TR 29821 bench.py --gpus 8 --skip-cpu --skip-parity --skip-exposed --e2e-sweep xpull=1 xstreams=1 xstreams=4 xgroups=1 xpull=1,h2d_sm=16 > gpurun_out/r2q_bench_reddit_n8.json 2> gpurun_out/r2q_bench_reddit_n8.err; echo "reddit rc=$? t=$(( $(date +%s) - t0 ))"
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/r2q_bench_reddit_n8.json"))
    print("value %.3f synced %.3f e2e %.3f" % (d["value"], d["synced_epoch"]["value"], d["e2e"]["value"]))
    print(json.dumps(d["e2e"].get("diag"), indent=1))
except Exception as ex:
    print("no line:", ex)
PY
tail -n 3 gpurun_out/r2q_bench_reddit_n8.err | cut -c1-400
timeout 120 python -m pytest tests/test_cpp_surface.py -m gpu -q -rs --timeout=110 -p no:cacheprovider -k "two_gpu_surface" > gpurun_out/r2q_pytest_surface_n2.log 2>&1; echo "pytest rc=$? t=$(( $(date +%s) - t0 ))"
tail -n 6 gpurun_out/r2q_pytest_surface_n2.log | cut -c1-300
