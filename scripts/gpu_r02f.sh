#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/r02f_pytest.log 2>&1
tail -3 gpurun_out/r02f_pytest.log; grep -E "^\.?(spheres1024|csg256|proctex|mixed4096) " gpurun_out/r02f_pytest.log | cut -c1-600
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r02f_bench.json 2> gpurun_out/r02f_bench.err
timeout 600 python bench.py --steps 10 --warmup 3 --parts 8 --no-cpu-baseline --no-parity > gpurun_out/r02f_bench_parts8.json 2>> gpurun_out/r02f_bench.err
python - <<'PY'
import json
for f in ("r02f_bench.json","r02f_bench_parts8.json"):
    b=json.load(open("gpurun_out/"+f)); print(f, round(b["ms_per_step"],3), round(b["e2e"]["frame_ms"],3), b["e2e"]["phase_ms"], round(b["e2e_pageable"]["frame_ms"],3), b["e2e_pageable"]["phase_ms"], {k:round(v,3) for k,v in b["kernel_ms_per_frame"].items()}, b["roofline"]["kernel"], b["roofline"]["frac"]); print(b.get("parity")); 
PY
for c in csg256 proctex mixed4096; do timeout 900 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline >> gpurun_out/r02f_configs.jsonl 2>> gpurun_out/r02f_configs.err; done
python - <<'PY'
import json
for l in open("gpurun_out/r02f_configs.jsonl"):
    b=json.loads(l); print(b["config"]["workload"][:40], round(b["ms_per_step"],3), {k:round(v,3) for k,v in b["kernel_ms_per_frame"].items()}); print("   ", b.get("parity"))
PY
