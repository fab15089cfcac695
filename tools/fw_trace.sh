#!/bin/bash
O=gpurun_out; mkdir -p $O
CMBPO_LIB=build/libcmbpo_b200_trace.so CMBPO_TRACE_OUT=$O/fw_trace.bin timeout 300 python tools/wide_bench.py --widths 512 --rows ${ROWS:-100000} --layers ${LAYERS:-4} --reps 1 2>&1 | tail -1
ls -la $O/fw_trace.bin
