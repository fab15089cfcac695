#!/bin/bash
# Runs every umma_probe case in its own process under a timeout; log -> gpurun_out/probe.log
mkdir -p gpurun_out
P=build/umma_probe
L=gpurun_out/probe.log
: > $L
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv >> $L 2>&1
run() { echo "--- $*" >> $L; timeout 60 $P "$@" >> $L 2>&1; echo "exit=$?" >> $L; }
run ss 1 208 48 0
run ss 1 208 48 1
run ss 1 64 208 0
run ts 1 208 208 0
run ts 1 208 208 1
run ts 1 64 208 0
run chain
run ss 2 208 48 0
run ss 2 208 208 0
run ts 2 224 208 0
run ts 2 208 208 0
run ts 2 64 208 0
run time ss 1 208 208 200
run time ts 1 208 208 200
run time ts 1 112 208 200
run time ts 1 64 208 200
run time ss 2 208 208 200
run time ts 2 224 208 200
run time ts 2 208 208 200
run time ts 1 256 208 200
run time ss 1 256 208 200
cat $L
