#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for env in "GMLRT_SORT_SHADOWS=0" "GMLRT_SORT_SHADOWS=1" "GMLRT_SORT_SHADOWS=1 GMLRT_KEY_K=6" "GMLRT_SORT_SHADOWS=1 GMLRT_KEY_K=8"; do
  echo "== $env" >> gpurun_out/r02g.log
  env $env timeout 900 python bench.py --config mixed4096 --steps 3 --warmup 3 --no-cpu-baseline --no-parity 2>> gpurun_out/r02g.err | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print(round(b['ms_per_step'],2), {k:round(v,2) for k,v in b['kernel_ms_per_frame'].items()}, b['roofline']['scan_counters_per_frame']['cull_tests_by_family'])" >> gpurun_out/r02g.log
done
for env in "GMLRT_SORT_SHADOWS=1" "GMLRT_KEY_K=6" "GMLRT_KEY_K=3"; do
  echo "== spheres1024 $env" >> gpurun_out/r02g.log
  env $env timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-parity 2>> gpurun_out/r02g.err | python -c "
import json,sys
b=json.loads(sys.stdin.read()); print(round(b['ms_per_step'],3), {k:round(v,3) for k,v in b['kernel_ms_per_frame'].items()}, b['roofline']['scan_counters_per_frame']['cull_tests_by_family'])" >> gpurun_out/r02g.log
done
cat gpurun_out/r02g.log
