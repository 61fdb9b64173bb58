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
    # variants: STRESS_VARIANTS="label:ENV=v,ENV=v label2:..." (default: one plain loop); STRESS_BUDGET_S caps the wall time;
    # STRESS_WARM=1 runs a multi-process bench first.  CUDA_LAUNCH_BLOCKING=1 is NOT a usable variant: one host thread
    # drives both engines of dist_gcn, and a blocking launch of the first one spins in its peer barrier forever.
    runs=${1:-20}
    B=mg-gcn_b200/cpp/bin/test_surface
    log=gpurun_out/${tag}_stress.log
    t0=$(date +%s); budget=${STRESS_BUDGET_S:-100000}
    : > $log
    nvidia-smi --query-gpu=index,name --format=csv,noheader | tr '\n' ';' >> $log; echo >> $log
    if [ "${STRESS_WARM:-0}" = 1 ] && [ "$N" -gt 1 ]; then
        timeout 300 bash -c "$(declare -f launch); N=$N; launch --steps 5 --warmup 3 --skip-parity --skip-cpu --skip-exposed" > gpurun_out/${tag}_stress_bench.json 2> gpurun_out/${tag}_stress_bench.err
        echo "multi-process job before the loop: rc=$? $(head -c 200 gpurun_out/${tag}_stress_bench.json)" >> $log
    fi
    for variant in ${STRESS_VARIANTS:-plain:A=1}; do
        label=${variant%%:*}; envs=$(echo "${variant#*:}" | tr ',' ' ')
        bad=0; done_=0
        for i in $(seq 1 $runs); do
            [ $(( $(date +%s) - t0 )) -ge $budget ] && break
            env MGGCN_BARRIER_TIMEOUT_MS=3000 $envs timeout 60 $B dist > gpurun_out/${tag}_stress_run.txt 2>&1; rc=$?
            done_=$((done_ + 1))
            [ $rc -ne 0 ] && bad=$((bad + 1))
            echo "--- $label run $i rc=$rc t=$(( $(date +%s) - t0 ))s" >> $log
            grep -E "first losses|checksum|layer stack|FAILED|mggcn failure" gpurun_out/${tag}_stress_run.txt | cut -c1-160 >> $log
        done
        echo "$label: $((done_ - bad)) of $done_ runs green" | tee -a $log
    done ;;
*) echo "usage: $0 suite|bench|stress [tag] [...]"; exit 2 ;;
esac
