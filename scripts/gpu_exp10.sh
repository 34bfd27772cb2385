#!/bin/bash
# experiment: how k_cull<GENERAL> scales with the number of rays (depth 2 vs 3, 4K vs 1080p vs 8K)
mkdir -p gpurun_out; : > gpurun_out/exp10.jsonl; : > gpurun_out/exp10.err
run() { echo "== $*"; timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline "$@" 2>> gpurun_out/exp10.err | tail -1 >> gpurun_out/exp10.jsonl; }
run --depth 2
run --depth 3
run --depth 4
run --depth 3 --width 1920 --height 1080
run --depth 3 --width 7680 --height 4320
python - <<'PY'
import json
for l in open("gpurun_out/exp10.jsonl"):
    try: d = json.loads(l)
    except Exception: print("bad line", l[:200]); continue
    s = d["roofline"]["scan_counters_per_frame"]; r = d["rays_per_frame"]
    print(d["config"]["workload"][:44], "ms", round(d["ms_per_step"], 3), "serial", round(d["serialised_frame_ms"], 3), {k: round(v, 3) for k, v in d["kernel_ms_per_frame"].items()}, "reflect", r["reflect"], "general tests", s["cull_tests_by_family"]["general"], "packet", s["packet_tests"])
PY
