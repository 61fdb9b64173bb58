# 2-GPU call: is the layer-stack mismatch a race?  three plain runs, one with blocking launches
mkdir -p gpurun_out
B=mg-gcn_b200/cpp/bin/test_surface
for k in 1 2 3; do timeout 60 $B dist > gpurun_out/r2r_dist_$k.log 2>&1; echo "run $k rc=$?"; grep -c "EXPECT failed" gpurun_out/r2r_dist_$k.log; grep "layer stack" gpurun_out/r2r_dist_$k.log | head -8; done
CUDA_LAUNCH_BLOCKING=1 timeout 90 $B dist > gpurun_out/r2r_dist_blocking.log 2>&1; echo "blocking rc=$?"; grep -c "EXPECT failed" gpurun_out/r2r_dist_blocking.log; grep "layer stack" gpurun_out/r2r_dist_blocking.log | head -8
