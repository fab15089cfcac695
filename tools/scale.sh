#!/bin/bash
# Scaling run of bench.py on N GPUs of one box: tools/scale.sh N [extra bench args]; output gpurun_out/scale_nN*.json
N=${1:-2}; shift
O=gpurun_out; mkdir -p $O
TAG=$(echo -n "$*" | tr -c 'a-zA-Z0-9' '_')
if [ "$N" = "1" ]; then CMD=(python bench.py --gpus 1)
else CMD=(python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N); fi
timeout 900 "${CMD[@]}" --steps 20 --warmup 3 --no-cpu "$@" > $O/scale_n${N}${TAG}.json 2> $O/scale_n${N}${TAG}.err; echo "rc=$?"
python - $O/scale_n${N}${TAG}.json <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(f"N={d['n_gpus']} {d['config']['workload']} value {d['value']/1e6:.1f} M/s  ms/step {d['ms_per_step']:.3f}  e2e {d['e2e']['value']/1e6:.1f} M/s  "
          f"K1 {d['roofline']['achieved']:.0f} TF ({d['roofline']['frac_of_burst']:.3f} burst) share {d['roofline']['kernel_share_of_step']:.2f}  {d['details']['parallelism']}")
except Exception as e:
    print("no json line:", e)
PY
tail -3 $O/scale_n${N}${TAG}.err
