#!/bin/bash
# Round evidence run on the GPU box: GPU parity suite, bench (both arms, every workload), forward microbench + width
# sweep, ncu launch list of the bench command and full ncu captures of the dominant kernels.
# Usage (from the repo root, after `make`): gpurun --timeout 1800 -- 'bash tools/gpu_evidence.sh r01'
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
timeout 900 python tools/wide_bench.py --widths 200,256,512,1024 --rows 65536,262144,1000000 > $O/${TAG}_sweep.log 2>&1; tail -14 $O/${TAG}_sweep.log
python tools/pack_bench.py > $O/${TAG}_pack_bench.log 2>&1; cat $O/${TAG}_pack_bench.log
b() { timeout 600 python bench.py --workload $1 --steps $2 --warmup 3 --no-cpu > $O/${TAG}_$3.json 2> $O/${TAG}_$3.err; echo "$3 rc=$?"; }
b halfcheetah_schedule_L10_1M 5 bench_halfcheetah
b antsafe_step_1M 10 bench_step1M
b antsafe_shipped_2x512_uncertainty_H30_100k 10 bench_wide512
b antsafe_step_1M_4x512 5 bench_step1M_4x512
b antsafe_step_256k_4x1024 5 bench_step256k_4x1024
b humanoidsafe_iv_gae_H12_125k 5 bench_iv_gae
# launch list (cold-cache, serialised: shares only)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-graph > $O/${TAG}_ncu_launches.log 2>&1; echo "ncu list rc=$?"
# full captures: fused forward and post kernels at 1M rows, one wide layer launch, the 3-D post kernel, the pack kernel
cap() { timeout 900 ncu --set full --clock-control none --import-source on -k regex:$1 -s $2 -c 1 -o $O/${TAG}_$3 "${@:4}" > $O/${TAG}_ncu_$3.log 2>&1; echo "ncu $3 rc=$?"; }
cap ensemble_fwd 3 fwd_1M python tools/fwd_bench.py 1000000 3
cap step_post_kernel 3 post_1M python tools/fwd_bench.py 1000000 3
cap wide_layer 6 wide_512 python tools/wide_bench.py --widths 512 --rows 262144 --reps 2
cap step_post3d 2 post3d python bench.py --workload humanoidsafe_iv_gae_H12_125k --steps 1 --warmup 3 --no-cpu --no-graph
cap k_pack_write 1 pack python tools/pack_bench.py
ls -la $O
