#!/bin/bash
# both arms of bench.py at N GPUs (the driver's launch line); JSON lines land in gpurun_out/
N=$1; shift
PORT=$((29500 + N))
run() { if [ "$N" = 1 ]; then python bench.py --gpus 1 "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@"; fi; }
run --impl reference --steps 5 --warmup 3 > gpurun_out/scale_ref_n$N.json 2> gpurun_out/scale_ref_n$N.err
run --steps 10 --warmup 3 "$@" > gpurun_out/scale_ours_n$N.json 2> gpurun_out/scale_ours_n$N.err
tail -c 1500 gpurun_out/scale_ref_n$N.json; echo; tail -c 2500 gpurun_out/scale_ours_n$N.json
