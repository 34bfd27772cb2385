#!/bin/bash
# plain run, then one --set full capture of the kernels matching $1 (regex), skipping $2 matches, capturing $3
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 280 $CMD > gpurun_out/plain2.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"$1" -s ${2:-0} -c ${3:-8} -o gpurun_out/cap_full -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -n 2 gpurun_out/plain2.log | cut -c1-300
