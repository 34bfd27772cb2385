#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/r02e_pytest.log 2>&1
tail -5 gpurun_out/r02e_pytest.log; grep -E "^(spheres1024|csg256|proctex|mixed4096) " gpurun_out/r02e_pytest.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r02e_bench.json 2> gpurun_out/r02e_bench.err
python - <<'PY'
import json
b=json.load(open("gpurun_out/r02e_bench.json")); print(round(b["ms_per_step"],3), b["e2e"]["frame_ms"], b["e2e"]["phase_ms"], b["e2e_pageable"]["frame_ms"], b["e2e_pageable"]["phase_ms"], {k:round(v,3) for k,v in b["kernel_ms_per_frame"].items()}); print(b.get("parity")); print(b.get("cpu_baseline"))
PY
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-parity"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02e_launches.csv $CMD > gpurun_out/r02e_ncu_list.log 2>&1
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"k_cull|k_shade|k_primary_points|k_resolve|k_exact" -s 70 -c 36 -o gpurun_out/r02e_full -f $CMD > gpurun_out/r02e_ncu_full.log 2>&1; echo "ncu full rc=$?"
