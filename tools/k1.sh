#!/bin/bash
# K1 iteration: parity of predict/step (small + at size), forward timing, in-kernel timeline
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_predict.py tests/test_gpu_step.py tests/test_gpu_at_size.py -m gpu -q -x > $O/k1_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/k1_pytest.log
for B in 100000 1000000; do timeout 300 python tools/fwd_bench.py $B 10 antsafe_step_1M 2>&1 | tail -1; done
timeout 300 python tools/fwd_bench.py 1000000 10 halfcheetah_schedule_L10_1M 2>&1 | tail -1
CMBPO_LIB=build/libcmbpo_b200_trace.so CMBPO_TRACE_OUT=$O/trace.bin timeout 300 python tools/fwd_bench.py 1000000 1 antsafe_step_1M > $O/trace_run.log 2>&1; echo "trace rc=$?"
python tools/trace_report.py $O/trace.bin 300 420 > $O/trace_report.txt 2>&1; tail -22 $O/trace_report.txt
