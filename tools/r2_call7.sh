# 2-GPU call: layer-stack diagnostics, grouped-exchange tests on the new build, papers100M-shape path at 2 % scale
mkdir -p gpurun_out
mg-gcn_b200/cpp/bin/test_surface dist > gpurun_out/r2g_surface_dist.log 2>&1; echo "test_surface dist rc=$?"
tail -n 12 gpurun_out/r2g_surface_dist.log
python -m pytest tests/test_gpu_engine.py -m gpu -q -rs --timeout=900 -p no:cacheprovider -k "two_gpu and (grouped or sharded or uneven)" > gpurun_out/r2g_pytest_n2.log 2>&1; echo "pytest n2 rc=$?"
tail -n 3 gpurun_out/r2g_pytest_n2.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29731"
timeout 400 $TR bench.py --gpus 2 --shape papers100M --scale 0.02 --comm p2p-staged --steps 3 > gpurun_out/r2g_papers_s002_n2.json 2> gpurun_out/r2g_papers_s002_n2.err; echo "papers100M x0.02 rc=$?"
tail -n 4 gpurun_out/r2g_papers_s002_n2.err
head -c 1500 gpurun_out/r2g_papers_s002_n2.json; echo
python bench.py --impl reference --gpus 8 --shape papers100M > gpurun_out/r2g_ref_papers100M_n8.json 2> gpurun_out/r2g_ref_papers100M_n8.err; echo "reference papers100M rc=$?"
head -c 600 gpurun_out/r2g_ref_papers100M_n8.json; echo
