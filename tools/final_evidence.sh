#!/bin/bash
# refresh of the evidence that changed with the last kernels: GPU suite, default bench (both arms), launch list
TAG=${1:-r01}; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest_gpu.log; tail -4 $O/${TAG}_pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"; cut -c1-400 $O/${TAG}_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err; echo "ref rc=$?"
b() { timeout 600 python bench.py --workload $1 --steps $2 --warmup 3 --no-cpu > $O/${TAG}_$3.json 2> $O/${TAG}_$3.err; echo "$3 rc=$?"; }
b halfcheetah_schedule_L10_1M 5 bench_halfcheetah
b antsafe_step_1M 10 bench_step1M
b antsafe_shipped_2x512_uncertainty_H30_100k 10 bench_wide512
b humanoidsafe_iv_gae_H12_125k 5 bench_iv_gae
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-graph > $O/${TAG}_ncu_launches.log 2>&1; echo "ncu list rc=$?"
python -c "import __graft_entry__ as g; g.smoke()"
