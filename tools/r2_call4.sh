# 8-GPU call A: world-4/8 parity tests, then reference/own bench pairs for reddit, products, proteins shapes
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2d_topo.txt 2>&1; (nproc; free -g | head -2; numactl -H 2>/dev/null | head -12) >> gpurun_out/r2d_topo.txt
timeout 600 python -m pytest tests/test_gpu_engine.py -m gpu -q -rs --timeout=300 -p no:cacheprovider -k "eight_gpu or four_gpu" > gpurun_out/r2d_pytest_n8.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2d_pytest_n8.log
tail -4 gpurun_out/r2d_pytest_n8.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29711"
for shape in reddit products proteins; do
  timeout 400 $TR bench.py --gpus 8 --shape $shape --impl reference > gpurun_out/r2d_ref_${shape}_n8.json 2> gpurun_out/r2d_ref_${shape}_n8.err; echo "ref $shape rc=$?"
  timeout 400 $TR bench.py --gpus 8 --shape $shape > gpurun_out/r2d_bench_${shape}_n8.json 2> gpurun_out/r2d_bench_${shape}_n8.err; echo "ours $shape rc=$?"
  head -c 300 gpurun_out/r2d_ref_${shape}_n8.json; echo; head -c 300 gpurun_out/r2d_bench_${shape}_n8.json; echo
  grep -o '"parity".*"ok": [a-z]*}' gpurun_out/r2d_bench_${shape}_n8.json | tail -c 200; echo
done
