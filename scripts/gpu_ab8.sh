#!/bin/bash
# A/B on one box, same library: camera rays made by k_gen (GMLRT_FUSE_GEN=0) or by the camera-ray cull kernel (default)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
run() {
  for parts in 1 8; do
  timeout 300 python bench.py --steps 20 --warmup 3 --parts $parts --no-cpu-baseline --no-parity --no-other-configs 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print('$1 parts $parts', round(b['ms_per_step'],3), 'serial', round(b['serialised_frame_ms'],3), {k:round(v,3) for k,v in b['kernel_ms_per_frame'].items()})"
  done
}
GMLRT_SORT_PERM=0 run records
run perm
GMLRT_SORT_PERM=0 run records
run perm
for c in csg256 mixed4096; do
  for f in 0 1; do
  GMLRT_SORT_PERM=$f timeout 600 python bench.py --config $c --steps 4 --warmup 3 --no-cpu-baseline --no-parity --no-other-configs 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print('$c sort_perm=$f', round(b['ms_per_step'],3), {k:round(v,3) for k,v in b['kernel_ms_per_frame'].items()})"
  done
done
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/ab8_pytest.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/ab8_pytest.log | cut -c1-300
