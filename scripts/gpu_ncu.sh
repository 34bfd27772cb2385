#!/bin/bash
# ncu visit: launch list, then one --set full capture of the scan kernels (args: extra bench.py flags)
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline $*"
timeout 280 $CMD > gpurun_out/plain1.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 280 $CMD > gpurun_out/plain2.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'k_cull|k_exact' -s 0 -c 20 -o gpurun_out/scan_full -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -2 gpurun_out/plain2.log | cut -c1-400
