# 2-GPU call: surface tests with poisoned allocations (MGGCN_POISON=1): where is memory read that nobody wrote?
mkdir -p gpurun_out
B=mg-gcn_b200/cpp/bin/test_surface
MGGCN_POISON=1 timeout 60 $B gpu > gpurun_out/r2t_poison_gpu.log 2>&1; echo "poison gpu rc=$? fails=$(grep -c 'EXPECT failed' gpurun_out/r2t_poison_gpu.log)"; grep -v "^$" gpurun_out/r2t_poison_gpu.log | head -12 | cut -c1-220
MGGCN_POISON=1 timeout 60 $B dist > gpurun_out/r2t_poison_dist.log 2>&1; echo "poison dist rc=$? fails=$(grep -c 'EXPECT failed' gpurun_out/r2t_poison_dist.log)"; grep -v "^$" gpurun_out/r2t_poison_dist.log | head -30 | cut -c1-220
