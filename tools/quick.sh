#!/bin/bash
# quick GPU check after a kernel change: step/at-size/rollout parity + step timing at 1M rows
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_step.py tests/test_gpu_at_size.py tests/test_gpu_rollout.py tests/test_gpu_predict.py -m gpu -q -x > $O/quick_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/quick_pytest.log
timeout 300 python tools/fwd_bench.py 1000000 10 antsafe_step_1M 2>&1 | tail -1
timeout 300 python tools/fwd_bench.py 1000000 10 halfcheetah_schedule_L10_1M 2>&1 | tail -1
timeout 300 python tools/fwd_bench.py 100000 20 antsafe_step_1M 2>&1 | tail -1
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > $O/quick_bench.json 2> $O/quick_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open('gpurun_out/quick_bench.json').read().strip().splitlines()[-1])
print(f"{d['config']['workload']} value {d['value']/1e6:.1f} M/s  ms/step {d['ms_per_step']:.3f}  e2e {d['e2e']['value']/1e6:.1f} M/s  K1 {d['roofline']['achieved']:.0f} TF share {d['roofline']['kernel_share_of_step']:.3f}")
PY
