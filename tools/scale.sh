#!/bin/bash
# Scaling run of bench.py on N GPUs of one box: tools/scale.sh N [extra bench args]; output gpurun_out/scale_nN*.json
N=${1:-2}; shift
O=gpurun_out; mkdir -p $O
TAG=$(echo "$*" | tr -c 'a-zA-Z0-9\n' '_')
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@"; }
if [ "$N" = "1" ]; then run() { python bench.py --gpus 1 "$@"; }; fi
timeout 900 run --steps 20 --warmup 3 --no-cpu "$@" > $O/scale_n${N}${TAG}.json 2> $O/scale_n${N}${TAG}.err; echo "rc=$?"
tail -c 1800 $O/scale_n${N}${TAG}.json; tail -3 $O/scale_n${N}${TAG}.err
