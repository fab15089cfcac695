#!/bin/bash
# Round evidence run on the GPU box: GPU parity suite, bench (both arms), 1M-row forward microbench,
# ncu launch list of the bench command and one full ncu capture of the ensemble forward kernel.
# Usage (from the repo root, after `make`): gpurun --timeout 1500 -- 'bash tools/gpu_evidence.sh r01'
TAG=${1:-r01}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
tail -c 3000 $O/${TAG}_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err; echo "ref rc=$?"
for B in 65536 262144 1000000 4000000; do timeout 300 python tools/fwd_bench.py $B 10; done > $O/${TAG}_fwd_bench.log 2>&1
cat $O/${TAG}_fwd_bench.log
timeout 600 python bench.py --workload halfcheetah_schedule_L10_1M --steps 5 --warmup 3 --no-cpu > $O/${TAG}_bench_halfcheetah.json 2> $O/${TAG}_bench_halfcheetah.err; echo "hc rc=$?"
timeout 600 python bench.py --workload antsafe_step_1M --steps 10 --warmup 3 --no-cpu > $O/${TAG}_bench_step1M.json 2> $O/${TAG}_bench_step1M.err; echo "1M rc=$?"
# launch list (cold-cache, serialised: shares only)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-graph > $O/${TAG}_ncu_launches.log 2>&1; echo "ncu list rc=$?"
# one full capture of the forward kernel at 1M rows
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ensemble_fwd -s 3 -c 1 -o $O/${TAG}_fwd_1M \
    python tools/fwd_bench.py 1000000 3 > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:step_post -s 3 -c 1 -o $O/${TAG}_post_1M \
    python tools/fwd_bench.py 1000000 3 > $O/${TAG}_ncu_full_post.log 2>&1; echo "ncu post rc=$?"
ls -la $O
