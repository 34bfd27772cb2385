#!/bin/bash
# ncu visit: launch list of one bench step, then one --set full capture of a tile's worth of the main kernels.
# The .ncu-rep can exceed what gpurun copies back, so the CSV / text pages are exported on the box and the report
# itself is kept only when small.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline $*"
timeout 280 $CMD > gpurun_out/plain1.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 280 $CMD > gpurun_out/plain2.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'k_shade|k_cull|k_exact|k_resolve|k_gen|k_keysort|k_closest' -s 0 -c 30 -o /tmp/cap_full -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i /tmp/cap_full.ncu-rep --page raw --csv > gpurun_out/cap_raw.csv 2>/dev/null
ncu -i /tmp/cap_full.ncu-rep --page details > gpurun_out/cap_details.txt 2>/dev/null
ncu -i /tmp/cap_full.ncu-rep --page source --csv --kernel-name regex:k_shade --launch-count 1 > gpurun_out/cap_src_shade.csv 2>/dev/null
ncu -i /tmp/cap_full.ncu-rep --page source --csv --kernel-name regex:'k_cull<float, 0' --launch-count 1 > gpurun_out/cap_src_cull_general.csv 2>/dev/null
ls -la /tmp/cap_full.ncu-rep; [ $(stat -c %s /tmp/cap_full.ncu-rep) -lt 30000000 ] && cp /tmp/cap_full.ncu-rep gpurun_out/
du -sh gpurun_out
tail -n 2 gpurun_out/plain2.log | cut -c1-300
