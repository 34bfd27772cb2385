#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/r02k_pytest.log 2>&1
tail -8 gpurun_out/r02k_pytest.log | cut -c1-300; grep "^ties" gpurun_out/r02k_pytest.log
GMLRT_DEBUG_UPLOAD=1 python scripts/gpu_io_diag.py 2>&1 | grep -v "^Exception\|^Traceback\|^  File\|^Attr" | tail -7
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-parity"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/r02k_launches.csv $CMD > gpurun_out/r02k_ncu_list.log 2>&1
