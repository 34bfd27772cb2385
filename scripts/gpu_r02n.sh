#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02n_noalloc_alone.log 2>&1; tail -3 gpurun_out/r02n_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02n_pytest.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/r02n_pytest.log | cut -c1-200
timeout 600 python scripts/gpu_icfp_diag.py chess 2>&1 | tail -40
