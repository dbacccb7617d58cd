#!/bin/bash
# Launch list (per-launch device time, cold-cache and serialised: compare SHARES) of one bench.py run: tools/ncu_launches.sh <out-name> <bench args...>
set -u
mkdir -p gpurun_out
out=$1; shift
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file "gpurun_out/$out.csv" \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-secondary "$@" > "gpurun_out/$out.log" 2>&1
python tools/ncu_summary.py "gpurun_out/$out.csv" 40
