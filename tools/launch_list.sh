#!/bin/bash
# per-kernel times (ncu launch list, cold-cache & serialised: compare shares) for one eager rollout of a bench workload
W=${1:-halfcheetah_schedule_L10_1M}; SKIP=${2:-150}; CNT=${3:-120}
python bench.py --workload $W --steps 2 --warmup 3 --no-cpu --no-graph > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s $SKIP -c $CNT --csv --log-file gpurun_out/launches_$W.csv python bench.py --workload $W --steps 2 --warmup 3 --no-cpu --no-graph > gpurun_out/ncu.log 2>&1; echo rc=$?
