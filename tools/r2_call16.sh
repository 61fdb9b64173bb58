# 2-GPU call: L2 read peak, the end-to-end step when the input copy is the longer side (narrow hidden layers make the
# epoch shorter than the 283 MB copy), a subset of the 2-GPU tests on the current build
mkdir -p gpurun_out
t0=$(date +%s)
timeout 120 python tools/l2_peak.py --mb 16 32 48 64 96 > gpurun_out/r2p_l2_peak.jsonl 2> gpurun_out/r2p_l2_peak.err; echo "l2_peak rc=$? t=$(( $(date +%s) - t0 ))"; cat gpurun_out/r2p_l2_peak.jsonl | cut -c1-80
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
run() { tag=$1; shift; timeout 120 $TR 29811 bench.py --gpus 2 --skip-parity --skip-cpu --skip-exposed "$@" > gpurun_out/r2p_${tag}.json 2> gpurun_out/r2p_${tag}.err; echo "$tag rc=$? t=$(( $(date +%s) - t0 ))"; python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2p_${tag}.json"))
    print("  value %.3f synced %.3f e2e %.3f" % (d["value"], d["synced_epoch"]["value"], d["e2e"]["value"]), d["e2e"].get("diag"))
except Exception as ex:
    print("  no line:", ex)
PY
}
run h16 --hidden 16 16 16
run h16_x1 --hidden 16 16 16 --tune xstreams=1
run h16_staged --hidden 16 16 16 --comm p2p-staged
timeout 300 python -m pytest tests/test_gpu_engine.py tests/test_cpp_surface.py -m gpu -q -rs --timeout=280 -p no:cacheprovider -k "two_gpu and (grouped or bcast or sharded or surface)" > gpurun_out/r2p_pytest_n2.log 2>&1; echo "pytest rc=$? t=$(( $(date +%s) - t0 ))"
tail -n 4 gpurun_out/r2p_pytest_n2.log
