#!/bin/bash
# experiment: k_shade software pipeline, camera ray recomputed on demand, K=4 default
mkdir -p gpurun_out; : > gpurun_out/exp7.jsonl; : > gpurun_out/exp7.err
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
run() { echo "== $ARGS $*"; env "$@" timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline $ARGS 2>> gpurun_out/exp7.err | tail -1 >> gpurun_out/exp7.jsonl; }
ARGS=""
run A=1
ARGS="--config proctex"
run A=1
ARGS="--config csg256"
run A=1
python - <<'PY'
import json
for l in open("gpurun_out/exp7.jsonl"):
    try: d = json.loads(l)
    except Exception: print("bad line", l[:200]); continue
    s = d["roofline"]["scan_counters_per_frame"]
    print(d["config"]["workload"][:40], "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["frame_ms"], 2), {k: round(v, 3) for k, v in d["kernel_ms_per_frame"].items()}, "cull", s["cull_tests_by_family"], "packet", s["packet_tests"], "exact", s["exact_tests"], s["exact_tests_shadow"])
PY
tail -5 gpurun_out/exp7.err
