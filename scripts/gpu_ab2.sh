#!/bin/bash
# A/B on one box: fused exact stage on / off (same library), and the previous build
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
run() {
  for parts in 1 8; do
  timeout 300 python bench.py --steps 20 --warmup 3 --parts $parts --no-cpu-baseline --no-parity 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print('$1 parts $parts', round(b['ms_per_step'],3), 'serial', round(b['serialised_frame_ms'],3), 'launches/frame', b['gpu_launches']//20, {k:round(v,3) for k,v in b['kernel_ms_per_frame'].items()})"
  done
}
GMLRT_LIB=$PWD/go-raytracer_b200/libgmlrt_base.so run base
GMLRT_FUSE_EXACT=0 run nofuse
run fuse
GMLRT_FUSE_EXACT=0 run nofuse
run fuse
for c in nothing; do
  for f in 0 1; do
  GMLRT_FUSE_EXACT=$f timeout 600 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline --no-parity 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print('$c fuse=$f', round(b['ms_per_step'],3), {k:round(v,3) for k,v in b['kernel_ms_per_frame'].items()})"
  done
done
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/ab2_pytest.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/ab2_pytest.log | cut -c1-300
