#!/bin/bash
# parity + every config, short
mkdir -p gpurun_out; : > gpurun_out/exp12.jsonl; : > gpurun_out/exp12.err
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
run() { echo "== $ARGS $*"; env "$@" timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline $ARGS 2>> gpurun_out/exp12.err | tail -1 >> gpurun_out/exp12.jsonl; }
for ARGS in "" "--parts 8" "--config proctex" "--config csg256" "--config mixed4096 --steps 3"; do run A=1; done
python - <<'PY'
import json
for l in open("gpurun_out/exp12.jsonl"):
    try: d = json.loads(l)
    except Exception: print("bad line", l[:200]); continue
    print(d["config"]["workload"][:34], d["config"]["workload"][-8:], "ms", round(d["ms_per_step"], 3), "serial", round(d["serialised_frame_ms"], 3), "e2e", round(d["e2e"]["frame_ms"], 2), {k: round(v, 2) for k, v in d["kernel_ms_per_frame"].items()})
PY
tail -3 gpurun_out/exp12.err
