#!/bin/bash
# timing experiments on the fused wide kernel: builds of the same library with one switch each (results of the NO* builds are wrong)
# usage (here, CPU): bash tools/fw_variants.sh build   |  (GPU box): bash tools/fw_variants.sh run
SRC=constrainedmbpo_b200/csrc
VARS="base:-DFW_GROUP=2 notanh:-DFW_EXP_NOTANH nofence:-DFW_EXP_NOFENCE nosts:-DFW_EXP_NOSTS noepi:-DFW_EXP_NOEPI halfw:-DFW_EXP_HALFW testpoll:-DFW_EXP_TESTPOLL plainwait:-DFW_EXP_PLAINWAIT group4:-DFW_GROUP=4 $FW_EXTRA_VARS"
if [ "$1" = build ]; then
  mkdir -p build/fwv
  for v in $VARS; do
    name=${v%%:*}; flag=${v#*:}
    nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC ${flag//,/ } -c $SRC/fused_wide.cu -o build/fwv/fused_wide_$name.o &
  done; wait
  for v in $VARS; do
    name=${v%%:*}
    nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/fwv/lib_$name.so build/ensemble_fwd.o build/wide_fwd.o build/fwv/fused_wide_$name.o build/step_post.o build/cmbpo_abi.o build/rollout_ops.o build/gae_ops.o build/archive_ops.o build/actor_ops.o -lcudart
  done
  ls -la build/fwv/*.so
else
  for v in $VARS; do
    name=${v%%:*}
    echo "== $name"
    CMBPO_LIB=build/fwv/lib_$name.so timeout 200 python tools/wide_bench.py --widths ${FW_W:-512} --rows 1000000 --layers 4 --reps 3 2>&1 | tail -1
    CMBPO_LIB=build/fwv/lib_$name.so timeout 200 python tools/wide_bench.py --widths ${FW_W:-512} --rows 1000000 --layers 2 --reps 3 2>&1 | tail -1
  done | tee gpurun_out/fw_variants.log
fi
