#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_wide.py tests/test_gpu_tf32.py tests/test_gpu_actor.py -m gpu -q -x > $O/fw_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/fw_pytest.log
for NSTR in 2 1; do
  echo "== CMBPO_FW_STREAMS=$NSTR"
  CMBPO_FW_STREAMS=$NSTR timeout 300 python tools/wide_bench.py --widths 256 --rows 25000,100000,1000000 --layers 4 --reps 3 2>&1 | tail -3
  CMBPO_FW_STREAMS=$NSTR timeout 300 python tools/wide_bench.py --widths 256 --rows 100000,1000000 --layers 2 --reps 3 2>&1 | tail -2
done | tee $O/fw_streams.log
