mkdir -p gpurun_out
python -m pytest tests -m gpu -q -rs --timeout=900 -p no:cacheprovider > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
tail -4 gpurun_out/r2b_pytest.log
python tools/l2_peak.py > gpurun_out/r2b_l2peak.jsonl 2>&1
python tools/spmm_bench.py --shape reddit --F 128 --cusparse --variants "auto:" "r01:blocks=1" "unsliced:blocks=1,slices=1" "persist:blocks=1,persist=1" "persistw2:blocks=1,persist=1,persist_waves=2" "tma128:blocks=1,slices=1,tma=1" "tma64:blocks=1,slices=2,tma=1" > gpurun_out/r2b_spmm_reddit.jsonl 2>&1
python tools/spmm_bench.py --shape reddit --F 44 --variants "auto:" "persist:persist=1" > gpurun_out/r2b_spmm_reddit_f44.jsonl 2>&1
python tools/spmm_bench.py --shape reddit --F 128 --blocks 8 --variants "np:persist=2" "p:persist=1" "p2:persist=1,persist_waves=2" "p4:persist=1,persist_waves=4" > gpurun_out/r2b_spmm_reddit_blk8.jsonl 2>&1
python tools/spmm_bench.py --shape products --F 256 --variants "unblocked:blocks=1" "auto:" "b6w32np:blocks=6,window=32,persist=2" "b6w32p2:blocks=6,window=32,persist=1,persist_waves=2" "b6w64:blocks=6,window=64" "b12w64:blocks=12,window=64" "b3w32:blocks=3,window=32" "b3w64:blocks=3,window=64" "b2w128:blocks=2,window=128" > gpurun_out/r2b_spmm_products.jsonl 2>&1
python tools/spmm_bench.py --shape proteins --F 128 --variants "auto:" "r01:blocks=1" "sl2:blocks=1,slices=2" "b3:blocks=3,window=128" > gpurun_out/r2b_spmm_proteins.jsonl 2>&1
tail -n 3 gpurun_out/r2b_spmm_products.jsonl
CMD1="python tools/spmm_bench.py --shape products --F 256 --variants b6w32:blocks=6,window=32"
$CMD1 > gpurun_out/r2b_plain1.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:spmm_kernel -s 5 -c 5 -o gpurun_out/r2b_prof_products $CMD1 > gpurun_out/r2b_ncu1.log 2>&1
CMD2="python tools/spmm_bench.py --shape reddit --F 128 --variants tma128:blocks=1,slices=1,tma=1"
$CMD2 > gpurun_out/r2b_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:spmm_tma -s 2 -c 1 -o gpurun_out/r2b_prof_tma $CMD2 > gpurun_out/r2b_ncu2.log 2>&1
CMD3="python tools/gemm_bench.py"
$CMD3 > gpurun_out/r2b_gemm_bench.jsonl 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 14 -c 3 -o gpurun_out/r2b_prof_gemm $CMD3 > gpurun_out/r2b_ncu3.log 2>&1
ls -la gpurun_out | tail -12
