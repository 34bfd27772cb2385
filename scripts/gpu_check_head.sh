#!/bin/bash
# what the driver runs at round end, on one box: GPU tests, smoke(), the default bench line and the reference arm
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/head_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/head_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/head_bench.json 2> gpurun_out/head_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
b=json.loads(open('gpurun_out/head_bench.json').read().strip().splitlines()[-1])
print({k:b.get(k) for k in ('value','ms_per_step','gpu_launches')}, b['e2e']['frame_ms'], b['e2e_new_scene']['frame_ms'], b['e2e_pageable']['frame_ms'], b['roofline']['frac'], b['roofline']['traffic'], b['parity']['within_1_255'], b['parity']['primary_flips'])
PY
