#!/bin/bash
# ncu --set full capture of the fused wide kernel (4x512 antsafe, 100k rows)
O=gpurun_out; mkdir -p $O
timeout 300 python tools/wide_bench.py --widths 512 --rows 100000 --layers 4 --reps 1 > $O/fw_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fused_wide -s 2 -c 1 -f -o $O/fw_4x512 python tools/wide_bench.py --widths 512 --rows 100000 --layers 4 --reps 1 > $O/fw_ncu.log 2>&1; echo "ncu rc=$?"
tail -3 $O/fw_ncu.log
