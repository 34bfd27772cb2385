#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/r02j_pytest.log 2>&1
tail -15 gpurun_out/r02j_pytest.log | cut -c1-300
GMLRT_DEBUG_UPLOAD=1 python scripts/gpu_io_diag.py 2>&1 | tail -9
for k in 3 5 6; do echo "parts8 k=$k"; GMLRT_KEY_K=$k timeout 600 python bench.py --steps 10 --warmup 3 --parts 8 --no-cpu-baseline --no-parity 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print(round(b['ms_per_step'],3), {k:round(v,3) for k,v in b['kernel_ms_per_frame'].items()})"; done
for k in 3 4; do echo "mixed parts8 k=$k"; GMLRT_KEY_K=$k timeout 600 python bench.py --config mixed4096 --steps 3 --warmup 3 --parts 8 --no-cpu-baseline --no-parity 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print(round(b['ms_per_step'],3), {k:round(v,3) for k,v in b['kernel_ms_per_frame'].items()})"; done
