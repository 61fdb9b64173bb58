# 2-GPU dev call: new surface tests, exchange split over streams, SM-driven H2D
mkdir -p gpurun_out
python -m pytest tests/test_cpp_surface.py -m gpu -q -rs --timeout=900 -p no:cacheprovider -k "surface" > gpurun_out/r2e_pytest_surface.log 2>&1; echo "pytest surface rc=$?"
tail -5 gpurun_out/r2e_pytest_surface.log
python -m pytest tests/test_gpu_engine.py -m gpu -q -rs --timeout=900 -p no:cacheprovider -k "two_gpu" > gpurun_out/r2e_pytest_n2.log 2>&1; echo "pytest n2 rc=$?"
tail -3 gpurun_out/r2e_pytest_n2.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29721"
run() { tag=$1; shift; timeout 300 $TR bench.py --gpus 2 --skip-parity --skip-cpu "$@" > gpurun_out/r2e_${tag}.json 2> gpurun_out/r2e_${tag}.err; echo "$tag rc=$?"; python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2e_${tag}.json")); x=d["exchange"]
    print("  value %.3f e2e %.3f copy_ms %.3f GBps %.0f exposed %.3f" % (d["value"], d["e2e"]["value"], x["copy_ms_per_epoch"], x["GBps"], x["exposed_ms"]))
except Exception as ex:
    print("  no line:", ex)
PY
}
run reddit_x1 --shape reddit --tune xstreams=1
run reddit_x2 --shape reddit --tune xstreams=2
run reddit_x4 --shape reddit --tune xstreams=4
run reddit_pull --shape reddit --tune xpull=1
run reddit_h2dsm --shape reddit --tune xstreams=2 h2d_sm=16
run products_x1 --shape products --tune xstreams=1
run products_x2 --shape products --tune xstreams=2
run products_x4 --shape products --tune xstreams=4
