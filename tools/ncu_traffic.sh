#!/bin/bash
# DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) of every kernel of one bench.py step:
#   tools/ncu_traffic.sh <out-name> <bench args...>      -> gpurun_out/<out-name>.csv, summarised by tools/ncu_traffic.py
set -u
mkdir -p gpurun_out
out=$1; shift
timeout 1500 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -c 6000 --csv \
  --log-file "gpurun_out/$out.csv" python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-secondary "$@" > "gpurun_out/$out.log" 2>&1
python tools/ncu_traffic.py "gpurun_out/$out.csv" 30
