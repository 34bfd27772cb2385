#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02p_pytest.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/r02p_pytest.log | cut -c1-300
GMLRT_DEBUG_UPLOAD=1 python scripts/gpu_io_diag.py 2>&1 | grep -v "^Exception\|^Traceback\|^  File\|^Attr" | tail -24
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --log-file gpurun_out/r02p_sanitizer_$tool.txt python scripts/gpu_sanitize.py 160 100 > gpurun_out/r02p_sanitizer_$tool.out 2>&1
  echo "$tool rc=$?"; tail -3 gpurun_out/r02p_sanitizer_$tool.txt
done
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r02p_bench.json 2> gpurun_out/r02p_bench.err
python - <<'PY'
import json
b=json.loads(open('gpurun_out/r02p_bench.json').read().strip().splitlines()[-1])
print({k:b.get(k) for k in ('value','ms_per_step','parity','gpu_launches')})
print(b['e2e']['frame_ms'], b['e2e']['phase_ms'], b['e2e_new_scene']['frame_ms'], b['e2e_new_scene']['phase_ms'], b['e2e_pageable']['frame_ms'])
PY
