#!/bin/bash
# fused wide kernel iteration: bit-exactness tests, then forward timing (fused only; pass LAYERED=1 for the A/B)
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_wide.py -m gpu -q -x -k "fused" > $O/fw_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/fw_pytest.log
for F in 1 ${LAYERED:+0}; do
  echo "== CMBPO_WIDE_FUSED=$F"
  CMBPO_WIDE_FUSED=$F timeout 300 python tools/wide_bench.py --widths 256,512 --rows 100000,1000000 --layers 4 --reps 3 2>&1 | tail -4
  CMBPO_WIDE_FUSED=$F timeout 300 python tools/wide_bench.py --widths 512 --rows 25000,100000,1000000 --layers 2 --reps 3 2>&1 | tail -3
done | tee $O/fw_wide_bench.log
