#!/bin/bash
# Fused wide kernel (fused_wide.cu) on one GPU: tools/fw.sh <mode>
#   check     bit-exactness / oracle tests of the wide kernels, then forward timing of 4x256, 4x512, 2x512 (LAYERED=1: also the
#             layer-at-a-time kernel, CMBPO_WIDE_FUSED=0)
#   streams   4x256 / 2x256 with two row-tile streams and with one (CMBPO_FW_STREAMS=1)
#   stages    weight-ring depth experiment (CMBPO_FW_STAGES)
#   variants  timing experiments: libraries built by `tools/fw_variants.sh build` with one feature switched off each (FW_W=width)
#   trace     in-kernel timeline of a -DCMBPO_TRACE build (`make trace`), printed by tools/fw_trace_report.py
#   ncu       ncu --set full capture of one 4x512 launch (100 k rows) -> gpurun_out/fw_4x512.ncu-rep
O=gpurun_out; mkdir -p $O
wb() { timeout 300 python tools/wide_bench.py "$@" 2>&1 | grep hidden; }
case "${1:-check}" in
check)
  timeout 900 python -m pytest tests/test_gpu_wide.py tests/test_gpu_tf32.py tests/test_gpu_actor.py -m gpu -q -x > $O/fw_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/fw_pytest.log
  for F in 1 ${LAYERED:+0}; do
    echo "== CMBPO_WIDE_FUSED=$F"
    CMBPO_WIDE_FUSED=$F wb --widths 256,512 --rows 100000,1000000 --layers 4 --reps 3
    CMBPO_WIDE_FUSED=$F wb --widths 512 --rows 25000,100000,1000000 --layers 2 --reps 3
  done | tee $O/fw_wide_bench.log ;;
streams)
  for NSTR in 2 1; do
    echo "== CMBPO_FW_STREAMS=$NSTR"
    CMBPO_FW_STREAMS=$NSTR wb --widths 256 --rows 25000,100000,1000000 --layers 4 --reps 3
    CMBPO_FW_STREAMS=$NSTR wb --widths 256 --rows 100000,1000000 --layers 2 --reps 3
  done | tee $O/fw_streams.log ;;
stages)
  for W in 384 512; do for S in 8 5 4 3; do
    echo -n "width $W stages<=$S: "; CMBPO_FW_STAGES=$S wb --widths $W --rows 1000000 --layers 4 --reps 3
  done; done | tee $O/fw_stages.log ;;
variants) bash tools/fw_variants.sh run ;;
trace)
  CMBPO_LIB=build/libcmbpo_b200_trace.so CMBPO_TRACE_OUT=$O/fw_trace.bin wb --widths ${FW_W:-512} --rows ${ROWS:-100000} --layers ${LAYERS:-4} --reps 1
  python tools/fw_trace_report.py $O/fw_trace.bin 300 420 | tail -60 ;;
ncu)
  wb --widths 512 --rows 100000 --layers 4 --reps 1 > $O/fw_plain.log && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:fused_wide -s 2 -c 1 -f -o $O/fw_4x512 python tools/wide_bench.py --widths 512 --rows 100000 --layers 4 --reps 1 > $O/fw_ncu.log 2>&1; echo "ncu rc=$?" ;;
*) echo "unknown mode $1"; exit 2 ;;
esac
