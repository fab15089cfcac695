#!/bin/bash
# weight-ring depth experiment: 4x384 (7 stages fit) and 4x512 (5 fit) with the ring capped
for W in 384 512; do for S in 8 5 4 3; do
  echo -n "width $W stages<=$S: "; CMBPO_FW_STAGES=$S timeout 200 python tools/wide_bench.py --widths $W --rows 1000000 --layers 4 --reps 3 2>&1 | tail -1
done; done | tee gpurun_out/fw_stages.log
