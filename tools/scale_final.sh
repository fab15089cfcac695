#!/bin/bash
# final weak-scaling line at N GPUs with the default flags (what the driver runs)
N=$1; O=gpurun_out; mkdir -p $O
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29800 bench.py --gpus $N --steps 20 --warmup 3 2>$O/scale_final_${N}.err | tail -1 > $O/r01_bench_n${N}.json
python -c "import json; d=json.loads(open('$O/r01_bench_n${N}.json').read()); print('N=$N value=%.1fM ms=%.3f e2e=%.1fM' % (d['value']/1e6, d['ms_per_step'], d['e2e']['value']/1e6))" || tail -5 $O/scale_final_${N}.err
