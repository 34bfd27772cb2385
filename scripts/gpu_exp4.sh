#!/bin/bash
# experiment: one rank's share of an 8-GPU frame under band / key-cell variants; upload phase breakdown; plane f32 robust test
mkdir -p gpurun_out; : > gpurun_out/exp4.jsonl; : > gpurun_out/exp4.err
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
run() { echo "== $*"; env "$@" timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline $ARGS 2>> gpurun_out/exp4.err | tail -1 >> gpurun_out/exp4.jsonl; }
ARGS="--parts 8"
run A=1
run GMLRT_BAND_ROWS=135
run GMLRT_BAND_ROWS=32
run GMLRT_KEY_K=2
run GMLRT_KEY_K=1
run GMLRT_CTA_PASS=0
ARGS="--config proctex"
run A=1
ARGS="--config csg256"
run A=1
ARGS="--config mixed4096 --steps 2"
run A=1
python - <<'PY'
import json
for l in open("gpurun_out/exp4.jsonl"):
    try: d = json.loads(l)
    except Exception: print("bad line", l[:200]); continue
    s = d["roofline"]["scan_counters_per_frame"]
    print(d["config"]["workload"][:40], "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["frame_ms"], 2), {k: round(v, 3) for k, v in d["kernel_ms_per_frame"].items()}, "cull", s["cull_tests_by_family"], "packet", s["packet_tests"], "exact", s["exact_tests"])
PY
GMLRT_DEBUG_UPLOAD=1 timeout 300 python scripts/gpu_upload.py 2>&1 | tail -12
tail -5 gpurun_out/exp4.err
