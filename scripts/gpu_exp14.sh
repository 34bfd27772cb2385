#!/bin/bash
# experiment: 128-thread scan CTAs (variant library built with -DGMLRT_SCAN_THREADS=128)
mkdir -p gpurun_out; : > gpurun_out/exp14.jsonl
cd go-raytracer_b200 && cp libgmlrt.so /tmp/libgmlrt_256.so && cp libgmlrt_128.so libgmlrt.so && cd ..
run() { timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline "$@" 2>> gpurun_out/exp14.err | tail -1 >> gpurun_out/exp14.jsonl; }
run
run --parts 8
run --config mixed4096 --steps 3
python - <<'PY'
import json
for l in open("gpurun_out/exp14.jsonl"):
    d = json.loads(l)
    print(d["config"]["workload"][:34], d["config"]["workload"][-8:], "ms", round(d["ms_per_step"], 3), "serial", round(d["serialised_frame_ms"], 3), {k: round(v, 3) for k, v in d["kernel_ms_per_frame"].items()})
PY
tail -3 gpurun_out/exp14.err
