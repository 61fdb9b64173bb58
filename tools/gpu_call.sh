#!/bin/bash
# One runner for the GPU-box calls of a round (replaces the one-shot call scripts of round 2).
#   tools/gpu_call.sh suite  [tag]          whole `-m gpu` suite, log to gpurun_out/<tag>_pytest.log
#   tools/gpu_call.sh bench  [tag] [args]   one bench.py line (N from $NGPU, default 1) -> gpurun_out/<tag>_bench.json
#   tools/gpu_call.sh stress [tag] [runs]   the in-process multi-GPU layer stack (test_surface dist) `runs` times, right
#                                           after a multi-process job on the same box (the condition of the one recorded
#                                           mismatch, profiles/r02_surface_dist_8gpu_box_first_run.log); on a failure the
#                                           loop is repeated with CUDA_LAUNCH_BLOCKING=1 and with poisoned allocations
# Everything lands in gpurun_out/ (merged back by gpurun); nothing here reads /root/reference.
set -u
mode=${1:-suite}; tag=${2:-call}; shift 2 2>/dev/null || true
mkdir -p gpurun_out
N=${NGPU:-1}
launch() { if [ "$N" = 1 ]; then python bench.py --gpus 1 "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py --gpus "$N" "$@"; fi; }
case "$mode" in
suite)
    timeout 900 python -m pytest tests -m gpu -q -rs -p no:cacheprovider "$@" > gpurun_out/${tag}_pytest.log 2>&1
    echo "suite rc=$?"; tail -n 6 gpurun_out/${tag}_pytest.log ;;
bench)
    timeout 900 bash -c "$(declare -f launch); N=$N; launch $*" > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
    echo "bench rc=$?"; tail -c 3000 gpurun_out/${tag}_bench.json ;;
stress)
    runs=${1:-25}
    B=mg-gcn_b200/cpp/bin/test_surface
    log=gpurun_out/${tag}_stress.log
    : > $log
    nvidia-smi --query-gpu=index,name --format=csv,noheader >> $log
    if [ "$N" -gt 1 ]; then
        timeout 300 bash -c "$(declare -f launch); N=$N; launch --steps 5 --warmup 3 --skip-parity --skip-cpu --skip-exposed" > gpurun_out/${tag}_stress_bench.json 2> gpurun_out/${tag}_stress_bench.err
        echo "multi-process job before the loop: rc=$? $(head -c 200 gpurun_out/${tag}_stress_bench.json)" >> $log
    fi
    loop() {  # $1 = label, rest = env assignments
        local label=$1; shift; local bad=0
        for i in $(seq 1 $runs); do
            env MGGCN_BARRIER_TIMEOUT_MS=3000 "$@" timeout 120 $B dist > gpurun_out/${tag}_stress_run.txt 2>&1; rc=$?
            if [ $rc -ne 0 ]; then bad=$((bad + 1)); echo "--- $label run $i rc=$rc" >> $log; grep -v '^$' gpurun_out/${tag}_stress_run.txt | head -n 30 | cut -c1-200 >> $log; fi
        done
        echo "$label: $((runs - bad)) of $runs runs green" | tee -a $log
        return $bad
    }
    loop "plain" A=1
    if [ $? -ne 0 ]; then
        loop "CUDA_LAUNCH_BLOCKING=1" CUDA_LAUNCH_BLOCKING=1
        loop "MGGCN_POISON=1" MGGCN_POISON=1
        loop "two visible devices" CUDA_VISIBLE_DEVICES=0,1
    fi ;;
*) echo "usage: $0 suite|bench|stress [tag] [...]"; exit 2 ;;
esac
