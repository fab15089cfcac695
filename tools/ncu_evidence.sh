#!/bin/bash
# ncu evidence for the default workload (one GPU): launch list of an eager rollout + full captures of K1 / K2 / k_store at 1M rows
O=gpurun_out; mkdir -p $O
python bench.py --steps 2 --warmup 3 --no-cpu --no-graph > $O/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 150 -c 120 --csv --log-file $O/r02_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-graph > $O/ncu_list.log 2>&1; echo "list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"ensemble_fwd|step_post_kernel|k_store" -s 25 -c 3 -o $O/r02_hot python bench.py --steps 1 --warmup 3 --no-cpu --no-graph > $O/ncu_full.log 2>&1; echo "full rc=$?"
python tools/fwd_bench.py 1000000 3 antsafe_step_1M > $O/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"ensemble_fwd|step_post_kernel" -s 6 -c 2 -o $O/r02_antsafe_1M python tools/fwd_bench.py 1000000 3 antsafe_step_1M > $O/ncu_full2.log 2>&1; echo "full2 rc=$?"
