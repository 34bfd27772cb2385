#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/tests_pytest.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/tests_pytest.log | cut -c1-300
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-other-configs 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print('cfg2', round(b['ms_per_step'],3), 'e2e', round(b['e2e']['frame_ms'],3), 'pageable', round(b['e2e_pageable']['frame_ms'],3), 'new scene', round(b['e2e_new_scene']['frame_ms'],3), b['parity']); print({k:(round(v['frac'],3) if v.get('frac') else None) for k,v in b['roofline']['by_class'].items()})"
