#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for v in "524288 6" "2097152 6" "2097152 8"; do
  set -- $v
  GMLRT_KEY_BINS=$1 GMLRT_KEY_K=$2 timeout 600 python bench.py --config mixed4096 --steps 4 --warmup 3 --no-cpu-baseline 2>gpurun_out/ab4_err.txt | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print('mixed4096 bins $1 k $2', round(b['ms_per_step'],3), 'new-scene upload', round(b['e2e_new_scene']['phase_ms']['upload'],2), {k:round(v,3) for k,v in b['kernel_ms_per_frame'].items()}); print(b['parity'])"
done
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-parity --no-other-configs 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print('cfg2', round(b['ms_per_step'],3), {k:round(v,3) for k,v in b['kernel_ms_per_frame'].items()})"
timeout 900 python -m pytest tests -m gpu -q -x -k "not baseline_size" > gpurun_out/ab4_pytest.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/ab4_pytest.log | cut -c1-300
