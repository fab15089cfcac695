#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x > $O/r02_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/r02_pytest_gpu.log
b() { timeout 900 python bench.py --workload $1 --steps $2 --warmup 3 ${4:-} > $O/r02_$3.json 2> $O/r02_$3.err; echo "$3 rc=$?"; }
b antsafe_shipped_2x512_uncertainty_H30_100k 30 bench_wide512 --no-cpu
b antsafe_step_1M_4x512 10 bench_step1M_4x512 --no-cpu
python - <<'PY'
import json,glob
for f in ['gpurun_out/r02_bench_wide512.json','gpurun_out/r02_bench_step1M_4x512.json']:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1]); r = d['roofline']
        print(f"{d['config']['workload']:44s} {d['value']/1e6:9.2f} M/s  e2e {d['e2e']['value']/1e6:9.2f}  {r['kernel']} {r['achieved']:7.1f} TF frac {r['frac']:.3f} share {r['kernel_share_of_step']:.3f}")
    except Exception as e:
        print(f, 'no line', e)
PY
