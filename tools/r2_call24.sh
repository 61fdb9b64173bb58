# 1-GPU call: input-buffer flip on single-GPU engines (two captured graphs): the prefetch test and the end-to-end step
mkdir -p gpurun_out
timeout 40 python -m pytest tests/test_gpu_engine.py -m gpu -q --timeout=35 -p no:cacheprovider -k "prefetched or graph_replay" > gpurun_out/r2x_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r2x_pytest.log | cut -c1-300
timeout 45 python bench.py --skip-parity --skip-cpu --steps 10 > gpurun_out/r2x_bench.json 2> gpurun_out/r2x_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2x_bench.json')); print('value %.3f synced %.3f e2e %.3f clocks %s' % (d['value'], d['synced_epoch']['value'], d['e2e']['value'], d['clocks']['sm_mhz']))"
