#!/bin/bash
# Round evidence run on one GPU: full GPU suite, bench lines of every BASELINE config (both arms for the default),
# tolerance report, width/batch sweep, launch list and full ncu captures of the dominant kernels.
# Usage: gpurun --timeout 2400 -- 'bash tools/gpu_evidence.sh r02'
TAG=${1:-r02}
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest_gpu.log
timeout 600 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err; echo "ref rc=$?"
b() { timeout 900 python bench.py --workload $1 --steps $2 --warmup 3 ${4:-} > $O/${TAG}_$3.json 2> $O/${TAG}_$3.err; echo "$3 rc=$?"; }
b antsafe_predict_10k 200 bench_predict10k
b antsafe_uncertainty_H30_100k 100 bench_uncertainty100k
b humanoidsafe_iv_gae_H12_125k 20 bench_iv_gae125k --no-cpu
b humanoidsafe_iv_gae_H12_1M 5 bench_iv_gae1M --no-cpu
b antsafe_step_1M 50 bench_step1M --no-cpu
b antsafe_step_4M 20 bench_step4M --no-cpu
b antsafe_shipped_2x512_uncertainty_H30_100k 30 bench_wide512 --no-cpu
b antsafe_step_1M_4x512 10 bench_step1M_4x512 --no-cpu
b antsafe_step_256k_4x1024 10 bench_step256k_4x1024 --no-cpu
timeout 900 python bench.py --workload antsafe_step_1M --operand-dtype tf32 --steps 10 --warmup 3 --no-cpu > $O/${TAG}_bench_step1M_tf32.json 2> $O/${TAG}_bench_step1M_tf32.err; echo "tf32 rc=$?"
for f in $O/${TAG}_bench*.json; do python - $f <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r = d.get("roofline") or {}
    print(f"{sys.argv[1].split('/')[-1]:40s} {d['config']['workload']:44s} {d['value']/1e6:9.2f} M/s  e2e {d['e2e']['value']/1e6:9.2f}  K1 {r.get('achieved', 0):7.1f} TF frac {r.get('frac', 0):.3f}")
except Exception as e:
    print(sys.argv[1], "no line", e)
PY
done
timeout 900 python tools/tolerance_report.py > $O/${TAG}_tolerance_report.txt 2>&1; cat $O/${TAG}_tolerance_report.txt
for B in 65536 262144 1000000 4000000; do timeout 300 python tools/fwd_bench.py $B 10 antsafe_step_1M 2>&1 | tail -1; done | tee $O/${TAG}_fwd_bench.log
# width x batch sweep of BASELINE configs[4] (4 hidden layers; 200: ensemble_fwd_kernel, 256-512: fused_wide_kernel, 1024: wide_layer_kernel)
timeout 900 python tools/wide_bench.py --widths 200,256,384,512,1024 --rows 65536,262144,1000000 --layers 4 --reps 3 2>&1 | grep hidden | tee $O/${TAG}_sweep.log
timeout 300 python tools/wide_bench.py --widths 512 --rows 25000,100000,1000000 --layers 2 --reps 3 2>&1 | grep hidden | tee -a $O/${TAG}_sweep.log
