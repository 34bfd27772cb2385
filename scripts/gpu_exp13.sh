#!/bin/bash
mkdir -p gpurun_out; : > gpurun_out/exp13.jsonl
run() { echo "== $ARGS $*"; env "$@" timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline $ARGS 2>> gpurun_out/exp13.err | tail -1 >> gpurun_out/exp13.jsonl; }
ARGS=""
run A=1
run A=1
run GMLRT_PICKLESS=0
python - <<'PY'
import json
for l in open("gpurun_out/exp13.jsonl"):
    d = json.loads(l)
    print(d["config"]["workload"][:34], "ms", round(d["ms_per_step"], 3), "serial", round(d["serialised_frame_ms"], 3), {k: round(v, 3) for k, v in d["kernel_ms_per_frame"].items()})
PY
