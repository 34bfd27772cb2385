#!/bin/bash
# experiment: 32-ray packets (RPT 1) for levels with few rays per sort bin (GMLRT_THIN_BIN_RAYS)
mkdir -p gpurun_out; : > gpurun_out/exp11.jsonl; : > gpurun_out/exp11.err
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
run() { echo "== $ARGS $*"; env "$@" timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline $ARGS 2>> gpurun_out/exp11.err | tail -1 >> gpurun_out/exp11.jsonl; }
for ARGS in "" "--parts 8" "--config mixed4096 --steps 2"; do
  for t in 0 64 200 1000000000; do run GMLRT_THIN_BIN_RAYS=$t; done
done
python - <<'PY'
import json
for l in open("gpurun_out/exp11.jsonl"):
    try: d = json.loads(l)
    except Exception: print("bad line", l[:200]); continue
    s = d["roofline"]["scan_counters_per_frame"]
    print(d["config"]["workload"][:34], d["config"]["workload"][-8:], "ms", round(d["ms_per_step"], 3), "serial", round(d["serialised_frame_ms"], 3), "cull2", round(d["kernel_ms_per_frame"]["cull_secondary"], 3), "exact", round(d["kernel_ms_per_frame"]["exact"], 3), "general tests", s["cull_tests_by_family"]["general"], "packet", s["packet_tests"])
PY
tail -3 gpurun_out/exp11.err
