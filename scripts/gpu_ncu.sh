#!/bin/bash
# ncu visit: launch list of one bench step, then one --set full capture of a tile's worth of the main kernels
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline $*"
timeout 280 $CMD > gpurun_out/plain1.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 280 $CMD > gpurun_out/plain2.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'k_shade|k_cull|k_exact|k_resolve|k_gen|k_keysort|k_closest' -s 0 -c 40 -o gpurun_out/cap_full -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -n 2 gpurun_out/plain2.log | cut -c1-300
