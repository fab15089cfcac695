#!/bin/bash
# N-GPU comparison of the gather transports (N = $1, default 8)
N=${1:-8}; O=gpurun_out; mkdir -p $O
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 20 --warmup 3 "${@:2}" 2>$O/scale${N}_err.log | tail -1 | tee -a $O/scale${N}_lines.jsonl | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('N=$N ${*:2}', 'value=%.1fM ms=%.3f e2e=%.1fM' % (d['value']/1e6, d['ms_per_step'], d['e2e']['value']/1e6))" || tail -5 $O/scale${N}_err.log; }
run 29701 --gather none
run 29702 --gather full --transport p2p
run 29703 --gather full --nccl-max-ctas 8
run 29704 --gather full
