# 2-GPU call (final build): subset of the 2-GPU tests, reference-layer parity at P = 2, reddit bench line at N = 2
mkdir -p gpurun_out
t0=$(date +%s)
timeout 200 python -m pytest tests/test_gpu_engine.py tests/test_cpp_surface.py tests/test_ref_layers.py -m gpu -q -rs --timeout=190 -p no:cacheprovider -k "(two_gpu and (grouped or bcast or sharded or surface)) or ref_layers" > gpurun_out/r2w_pytest_n2.log 2>&1; echo "pytest rc=$? t=$(( $(date +%s) - t0 ))"
tail -n 6 gpurun_out/r2w_pytest_n2.log | cut -c1-200
ls gpurun_out/ref_layers_*P2.txt 2>/dev/null
timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29841 bench.py --gpus 2 --skip-cpu > gpurun_out/r2w_bench_reddit_n2.json 2> gpurun_out/r2w_bench_reddit_n2.err; echo "bench rc=$? t=$(( $(date +%s) - t0 ))"
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/r2w_bench_reddit_n2.json"))
    print("value %.3f e2e %.3f parity %s %s exposed %s" % (d["value"], d["e2e"]["value"], d["parity"]["ok"], d["parity"]["max_abs_diff"], d["exchange"].get("exposed_frac")))
except Exception as ex:
    print("no line:", ex)
PY
