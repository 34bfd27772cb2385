#!/bin/bash
# experiment: child-slot zeroing skipped (parity), direction cells per region (GMLRT_KEY_K) with one-tile frames
mkdir -p gpurun_out; : > gpurun_out/exp6.jsonl; : > gpurun_out/exp6.err
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
run() { echo "== $ARGS $*"; env "$@" timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline $ARGS 2>> gpurun_out/exp6.err | tail -1 >> gpurun_out/exp6.jsonl; }
ARGS=""
run A=1
run GMLRT_KEY_K=4
run GMLRT_KEY_K=5
run GMLRT_KEY_K=6
ARGS="--config mixed4096 --steps 2"
run GMLRT_KEY_K=4
run GMLRT_KEY_K=2
python - <<'PY'
import json
for l in open("gpurun_out/exp6.jsonl"):
    try: d = json.loads(l)
    except Exception: print("bad line", l[:200]); continue
    s = d["roofline"]["scan_counters_per_frame"]
    print(d["config"]["workload"][:40], "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["frame_ms"], 2), {k: round(v, 3) for k, v in d["kernel_ms_per_frame"].items()}, "cull", s["cull_tests_by_family"], "packet", s["packet_tests"], "exact", s["exact_tests"], s["exact_tests_shadow"])
PY
tail -5 gpurun_out/exp6.err
