#!/bin/bash
# final_evidence.sh + fresh full captures of the two post kernels (they changed last)
TAG=${1:-r01}; O=gpurun_out; mkdir -p $O
bash tools/final_evidence.sh $TAG
cap() { timeout 600 ncu --set full --clock-control none --import-source on -k regex:$1 -s $2 -c 1 -o $O/${TAG}_$3 "${@:4}" > $O/${TAG}_ncu_$3.log 2>&1; echo "ncu $3 rc=$?"; }
rm -f $O/${TAG}_post_1M.ncu-rep $O/${TAG}_post3d.ncu-rep
cap step_post_kernel 3 post_1M python tools/fwd_bench.py 1000000 3
cap step_post3d 2 post3d python bench.py --workload humanoidsafe_iv_gae_H12_125k --steps 1 --warmup 3 --no-cpu --no-graph
timeout 300 python tools/fwd_bench.py 1000000 3 > $O/${TAG}_fwd_bench.log 2>&1; tail -2 $O/${TAG}_fwd_bench.log
