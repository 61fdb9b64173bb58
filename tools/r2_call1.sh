mkdir -p gpurun_out
(nproc; free -g | head -2; nvidia-smi -L; lscpu | grep -E "Model name|NUMA node\(s\)|Socket") > gpurun_out/r2_box.txt 2>&1
python -m pytest tests -m gpu -q -rs --timeout=900 -p no:cacheprovider > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
python bench.py --impl reference > gpurun_out/r2a_ref_reddit_n1.json 2> gpurun_out/r2a_ref_reddit_n1.err
python bench.py > gpurun_out/r2a_bench_reddit_n1.json 2> gpurun_out/r2a_bench_reddit_n1.err
python bench.py --shape products --impl reference > gpurun_out/r2a_ref_products_n1.json 2> gpurun_out/r2a_ref_products_n1.err
python bench.py --shape products > gpurun_out/r2a_bench_products_n1.json 2> gpurun_out/r2a_bench_products_n1.err
python bench.py --shape proteins --impl reference > gpurun_out/r2a_ref_proteins_n1.json 2> gpurun_out/r2a_ref_proteins_n1.err
python bench.py --shape proteins > gpurun_out/r2a_bench_proteins_n1.json 2> gpurun_out/r2a_bench_proteins_n1.err
head -c 600 gpurun_out/r2a_bench_reddit_n1.json; echo; tail -3 gpurun_out/r2a_bench_products_n1.err
