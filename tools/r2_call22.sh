# 1-GPU call (final build): reddit bench line first (leaves device memory dirty), whole GPU suite, smoke, surface tests
# with poisoned allocations, ncu launch list of the bench command
mkdir -p gpurun_out
t0=$(date +%s)
timeout 400 python bench.py > gpurun_out/r2v_bench_reddit_n1.json 2> gpurun_out/r2v_bench_reddit_n1.err; echo "bench rc=$? t=$(( $(date +%s) - t0 ))"
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/r2v_bench_reddit_n1.json"))
    print("value %.3f synced %.3f e2e %.3f gemm %.3f spmm %.3f parity %s %s clocks %s l2 %s" % (d["value"], d["synced_epoch"]["value"], d["e2e"]["value"], d["breakdown_ms"]["gemm"], d["breakdown_ms"]["spmm"], d["parity"]["ok"], d["parity"]["max_abs_diff"], d["clocks"]["sm_mhz"], d["roofline"].get("l2")))
except Exception as ex:
    print("no line:", ex)
PY
timeout 600 python -m pytest tests -m gpu -q -rs --timeout=600 -p no:cacheprovider > gpurun_out/r2v_pytest.log 2>&1; echo "pytest rc=$? t=$(( $(date +%s) - t0 ))"
tail -n 11 gpurun_out/r2v_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2v_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 1 gpurun_out/r2v_smoke.log
MGGCN_POISON=1 timeout 60 mg-gcn_b200/cpp/bin/test_surface gpu > gpurun_out/r2v_poison_gpu.log 2>&1; echo "poison gpu rc=$?"; tail -n 2 gpurun_out/r2v_poison_gpu.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2v_launches.csv python bench.py --steps 2 --warmup 1 --skip-parity --skip-cpu > gpurun_out/r2v_ncu_launches.log 2>&1; echo "ncu launches rc=$? t=$(( $(date +%s) - t0 ))"
