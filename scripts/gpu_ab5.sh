#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for byc in 1 0; do
  echo "GMLRT_LISTS_BY_CLASS=$byc"
  GMLRT_LISTS_BY_CLASS=$byc GMLRT_DEBUG_UPLOAD=1 python scripts/gpu_io_diag.py 2>&1 | grep "lists dev\|^upload" | head -4 | cut -c1-330
  GMLRT_LISTS_BY_CLASS=$byc GMLRT_KEY_BINS=2097152 GMLRT_DEBUG_UPLOAD=1 timeout 600 python bench.py --config mixed4096 --steps 3 --warmup 3 --no-cpu-baseline --no-parity 2>gpurun_out/ab5_err.txt | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print('mixed4096 2M bins', round(b['ms_per_step'],3), 'new-scene upload', round(b['e2e_new_scene']['phase_ms']['upload'],2), 'cull_secondary', round(b['kernel_ms_per_frame']['cull_secondary'],2))"
  grep "lists dev" gpurun_out/ab5_err.txt | tail -1 | cut -c1-330
done
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print('cfg2', round(b['ms_per_step'],3), 'e2e', round(b['e2e']['frame_ms'],3), 'new-scene', round(b['e2e_new_scene']['frame_ms'],3), b['e2e_new_scene']['phase_ms'], b['parity'])"
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/ab5_pytest.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/ab5_pytest.log | cut -c1-300
