#!/bin/bash
# One `ncu --set full` capture of named kernels under a bench.py run (never a bench number): tools/ncu_kernel.sh <out-name> <kernel-regex> <count> <bench args...>
set -u
mkdir -p gpurun_out
out=$1; rx=$2; cnt=$3; shift 3
timeout 1200 ncu --set full --clock-control none --import-source on -k "regex:$rx" -c "$cnt" -f -o "gpurun_out/$out" \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-secondary "$@" > "gpurun_out/$out.log" 2>&1
tail -3 "gpurun_out/$out.log"
ls -la gpurun_out/$out.ncu-rep
