# 1-GPU call: sanity of the rebuilt surface binaries
timeout 25 mg-gcn_b200/cpp/bin/test_surface gpu 2>&1 | tail -n 2
