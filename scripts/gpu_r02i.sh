#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02i_pytest.log 2>&1
tail -3 gpurun_out/r02i_pytest.log
show() { python -c "
import json,sys
b=json.loads(open(sys.argv[1]).read().strip().split('\n')[-1]); print(sys.argv[1], round(b['ms_per_step'],3), 'e2e', round(b['e2e']['frame_ms'],3), {k:round(v,3) for k,v in b['e2e']['phase_ms'].items()}, 'pageable', round(b['e2e_pageable']['frame_ms'],3), {k:round(v,3) for k,v in b['kernel_ms_per_frame'].items()}, b['roofline']['scan_counters_per_frame']['cull_tests_by_family'])" $1; }
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-parity > gpurun_out/r02i_bench.json 2> gpurun_out/r02i_bench.err; show gpurun_out/r02i_bench.json
for p in 2 4 8; do timeout 600 python bench.py --steps 10 --warmup 3 --parts $p --no-cpu-baseline --no-parity > gpurun_out/r02i_bench_parts$p.json 2>> gpurun_out/r02i_bench.err; show gpurun_out/r02i_bench_parts$p.json; done
timeout 900 python bench.py --config mixed4096 --steps 3 --warmup 3 --no-cpu-baseline --no-parity > gpurun_out/r02i_mixed.json 2>> gpurun_out/r02i_bench.err; show gpurun_out/r02i_mixed.json
timeout 900 python bench.py --config mixed4096 --parts 8 --steps 3 --warmup 3 --no-cpu-baseline --no-parity > gpurun_out/r02i_mixed_parts8.json 2>> gpurun_out/r02i_bench.err; show gpurun_out/r02i_mixed_parts8.json
