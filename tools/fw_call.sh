#!/bin/bash
# fused wide kernel check: parity (bit-exact vs the layer-at-a-time kernel, oracle), then A/B timing
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_wide.py tests/test_gpu_tf32.py -m gpu -q -x > $O/fw_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $O/fw_pytest.log
for F in 1 0; do
  echo "== CMBPO_WIDE_FUSED=$F"
  CMBPO_WIDE_FUSED=$F timeout 300 python tools/wide_bench.py --widths 256,512 --rows 100000,1000000 --layers 4 --reps 3 2>&1 | tail -4
  CMBPO_WIDE_FUSED=$F timeout 300 python tools/wide_bench.py --widths 512 --rows 25000,100000,1000000 --layers 2 --reps 3 2>&1 | tail -3
done | tee $O/fw_wide_bench.log
for F in 1 0; do
  CMBPO_WIDE_FUSED=$F timeout 600 python bench.py --workload antsafe_shipped_2x512_uncertainty_H30_100k --steps 30 --warmup 3 --no-cpu > $O/fw_bench_wide512_$F.json 2> $O/fw_bench_wide512_$F.err; echo "bench fused=$F rc=$?"
done
python - <<'PY'
import json
for F in (1, 0):
    try:
        d = json.loads(open(f'gpurun_out/fw_bench_wide512_{F}.json').read().strip().splitlines()[-1])
        print(F, d['config']['workload'], f"value {d['value']/1e6:.1f} M/s  ms/step {d['ms_per_step']:.3f}  e2e {d['e2e']['value']/1e6:.1f}  K1 {d['roofline']['achieved']:.0f} TF share {d['roofline']['kernel_share_of_step']:.3f}")
    except Exception as e:
        print(F, 'no line', e)
PY
