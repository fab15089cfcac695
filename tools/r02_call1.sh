#!/bin/bash
# round-2 GPU call 1: full GPU suite (incl. the at-size tests), default bench, K1 timeline at 1M rows
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x --durations=15 > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -40 $O/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; tail -c 2500 $O/bench.json; tail -5 $O/bench.err
for B in 262144 1000000; do timeout 300 python tools/fwd_bench.py $B 10 antsafe_step_1M; done 2>&1 | tee $O/fwd_bench.log
timeout 300 python tools/fwd_bench.py 1000000 10 halfcheetah_schedule_L10_1M 2>&1 | tee -a $O/fwd_bench.log
CMBPO_LIB=build/libcmbpo_b200_trace.so CMBPO_TRACE_OUT=$O/trace.bin timeout 300 python tools/fwd_bench.py 1000000 1 antsafe_step_1M > $O/trace_run.log 2>&1; echo "trace rc=$?"
python tools/trace_report.py $O/trace.bin 0 700 > $O/trace_report.txt 2>&1; tail -30 $O/trace_report.txt
