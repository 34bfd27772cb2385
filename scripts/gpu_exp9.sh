#!/bin/bash
# experiment: shadow queries on a second stream (overlap on / off), parity, one rank's share
mkdir -p gpurun_out; : > gpurun_out/exp9.jsonl; : > gpurun_out/exp9.err
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
run() { echo "== $ARGS $*"; env "$@" timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline $ARGS 2>> gpurun_out/exp9.err | tail -1 >> gpurun_out/exp9.jsonl; }
ARGS=""
run A=1
run GMLRT_OVERLAP_SHADOWS=0
ARGS="--parts 8"
run A=1
run GMLRT_OVERLAP_SHADOWS=0
ARGS="--config mixed4096 --steps 3"
run A=1
ARGS="--config csg256"
run A=1
ARGS="--config proctex"
run A=1
python - <<'PY'
import json
for l in open("gpurun_out/exp9.jsonl"):
    try: d = json.loads(l)
    except Exception: print("bad line", l[:200]); continue
    print(d["config"]["workload"][:40], "ms", round(d["ms_per_step"], 3), "serial", round(d["serialised_frame_ms"], 3), "e2e", round(d["e2e"]["frame_ms"], 2), {k: round(v, 3) for k, v in d["kernel_ms_per_frame"].items()})
PY
tail -5 gpurun_out/exp9.err
