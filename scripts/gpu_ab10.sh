#!/bin/bash
# A/B on one box, same library: CSG objects met by camera rays in double (GMLRT_CSG_F32_FIRST=0) or float32-first with tie tracking
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for f in 0 1 0 1; do
  GMLRT_CSG_F32_FIRST=$f timeout 600 python bench.py --config csg256 --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print('csg256 f32_first=$f', round(b['ms_per_step'],3), {k:round(v,3) for k,v in b['kernel_ms_per_frame'].items()}); print('   ', b['parity'])"
done
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/ab10_pytest.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/ab10_pytest.log | cut -c1-300
