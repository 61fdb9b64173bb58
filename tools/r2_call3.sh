mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2c_topo.txt 2>&1; (nproc; free -g | head -2) >> gpurun_out/r2c_topo.txt
python -m pytest tests/test_gpu_engine.py tests/test_ref_layers.py tests/test_cpp_surface.py -m gpu -q -rs --timeout=900 -p no:cacheprovider > gpurun_out/r2c_pytest_n2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2c_pytest_n2.log
tail -6 gpurun_out/r2c_pytest_n2.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29701"
$TR bench.py --gpus 2 --impl reference > gpurun_out/r2c_ref_reddit_n2.json 2> gpurun_out/r2c_ref_reddit_n2.err
$TR bench.py --gpus 2 > gpurun_out/r2c_bench_reddit_n2.json 2> gpurun_out/r2c_bench_reddit_n2.err
head -c 1200 gpurun_out/r2c_bench_reddit_n2.json; tail -3 gpurun_out/r2c_bench_reddit_n2.err
