#!/bin/bash
O=gpurun_out
for V in ${VARIANTS:-0 1 2}; do
  CMBPO_K1_DBG=$V timeout 300 python tools/fwd_bench.py 1000000 10 antsafe_step_1M 2>&1 | tail -1
  CMBPO_K1_DBG=$V CMBPO_LIB=build/libcmbpo_b200_trace.so CMBPO_TRACE_OUT=$O/trace$V.bin timeout 300 python tools/fwd_bench.py 1000000 1 antsafe_step_1M > $O/trace_run.log 2>&1
  python tools/trace_report.py $O/trace$V.bin 400 520 > $O/trace_report$V.txt 2>&1
done
