#!/bin/bash
# bench.py under different values of one environment variable: gpu_env_sweep.sh VAR "extra bench flags" v1 v2 ...
VAR=$1; FLAGS=$2; shift; shift
mkdir -p gpurun_out
for v in "$@"; do
  env $VAR=$v timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline $FLAGS > gpurun_out/sweep_$v.json 2> gpurun_out/sweep_$v.err
  python - "$VAR" "$v" <<'PY'
import json, sys
try:
    d = json.load(open(f"gpurun_out/sweep_{sys.argv[2]}.json"))
    ws = d["roofline"]["scan_counters_per_frame"]
    print(sys.argv[1], sys.argv[2], "ms", round(d["ms_per_step"], 3), {k: round(v, 3) for k, v in d["kernel_ms_per_frame"].items()}, "general", ws["cull_tests_by_family"]["general"] // 1000000, "M packet", ws["packet_tests"] // 1000000, "M rays", sum(d["rays_per_frame"][k] for k in ("primary", "reflect", "refract", "shadow")) // 1000000, "M")
except Exception as e:
    print(sys.argv[1], sys.argv[2], "failed", e)
PY
done
