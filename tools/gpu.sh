#!/bin/bash
# Rebuild the library, then run a command on the GPU box: tools/gpu.sh <timeout-seconds> '<command>'
set -e
cd "$(dirname "$0")/.."
make >/dev/null 2>build/make.log || { tail -30 build/make.log; exit 1; }
T=$1; shift
exec /usr/local/graft/bin/gpurun --timeout "$T" -- "$@"
