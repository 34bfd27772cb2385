#!/bin/bash
# A/B on one box: previous build against the current one; the key-bin limit on 8K mixed4096
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
run() {
  for parts in 1 8; do
  timeout 300 python bench.py --steps 20 --warmup 3 --parts $parts --no-cpu-baseline --no-parity --no-other-configs 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print('$1 parts $parts', round(b['ms_per_step'],3), 'serial', round(b['serialised_frame_ms'],3), {k:round(v,3) for k,v in b['kernel_ms_per_frame'].items()})"
  done
}
GMLRT_LIB=$PWD/go-raytracer_b200/libgmlrt_base.so run base
run new
GMLRT_LIB=$PWD/go-raytracer_b200/libgmlrt_base.so run base
run new
for v in "524288 6" "1048576 6" "2097152 6" "1048576 8" "2097152 8"; do
  set -- $v
  GMLRT_DEBUG_UPLOAD=1 GMLRT_KEY_BINS=$1 GMLRT_KEY_K=$2 timeout 600 python bench.py --config mixed4096 --steps 4 --warmup 3 --no-cpu-baseline --no-parity --no-other-configs 2>gpurun_out/ab3_err.txt | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print('mixed4096 bins $1 k $2', round(b['ms_per_step'],3), 'new-scene upload', round(b['e2e_new_scene']['phase_ms']['upload'],2), {k:round(v,3) for k,v in b['kernel_ms_per_frame'].items()})"
  grep "lists dev" gpurun_out/ab3_err.txt | tail -1 | cut -c1-300
done
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/ab3_pytest.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/ab3_pytest.log | cut -c1-300
