#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02q_pytest.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/r02q_pytest.log | cut -c1-300
for p in 2 4 8; do
timeout 300 python bench.py --steps 20 --warmup 3 --parts $p --no-cpu-baseline --no-parity 2>/dev/null | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print('parts $p', round(b['ms_per_step'],3), 'serial', round(b['serialised_frame_ms'],3), {k:round(v,3) for k,v in b['kernel_ms_per_frame'].items()})"
done
timeout 300 python scripts/gpu_one_frame.py spheres1024 1 && \
timeout 900 ncu --set full --clock-control none --profile-from-start off -o gpurun_out/r02q_frame_full -f python scripts/gpu_one_frame.py spheres1024 1 > gpurun_out/r02q_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r02q_launches_parts8.csv python scripts/gpu_one_frame.py spheres1024 8 > gpurun_out/r02q_ncu_p8.log 2>&1
ls -la gpurun_out/*.ncu-rep
