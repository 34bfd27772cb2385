#!/bin/bash
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
for c in csg256 proctex; do
  for lib in base new; do
  if [ $lib = base ]; then export GMLRT_LIB=$PWD/go-raytracer_b200/libgmlrt_base.so; else unset GMLRT_LIB; fi
  timeout 600 python bench.py --config $c --steps 4 --warmup 3 --no-cpu-baseline --no-parity --no-other-configs 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print('$c $lib', round(b['ms_per_step'],3), {k:round(v,3) for k,v in b['kernel_ms_per_frame'].items()})"
  done
done
unset GMLRT_LIB
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/ab6_pytest.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/ab6_pytest.log | cut -c1-300
