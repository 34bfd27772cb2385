#!/bin/bash
# experiment: mixed4096 tile size / key cells sweep; spheres1024 one rank's share of an 8-GPU frame
mkdir -p gpurun_out; : > gpurun_out/exp1.jsonl
run() { echo "== $*"; env "$@" timeout 300 python bench.py --config mixed4096 --steps 2 --warmup 3 --no-cpu-baseline 2>> gpurun_out/exp1.err | tail -1 >> gpurun_out/exp1.jsonl; }
run GMLRT_POOL_MB=98304 GMLRT_TILE_SAMPLES=67108864
run GMLRT_POOL_MB=98304 GMLRT_TILE_SAMPLES=67108864 GMLRT_KEY_K=2
run GMLRT_POOL_MB=98304 GMLRT_TILE_SAMPLES=67108864 GMLRT_KEY_K=1
run GMLRT_KEY_K=1
timeout 300 python bench.py --parts 8 --steps 10 --warmup 3 --no-cpu-baseline 2>> gpurun_out/exp1.err | tail -1 >> gpurun_out/exp1.jsonl
timeout 300 python bench.py --parts 2 --steps 10 --warmup 3 --no-cpu-baseline 2>> gpurun_out/exp1.err | tail -1 >> gpurun_out/exp1.jsonl
python - <<'PY'
import json
for l in open("gpurun_out/exp1.jsonl"):
    try: d = json.loads(l)
    except Exception: print("bad line", l[:200]); continue
    s = d["roofline"]["scan_counters_per_frame"]
    print(d["config"]["workload"][:60], "ms", round(d["ms_per_step"], 2), {k: round(v, 2) for k, v in d["kernel_ms_per_frame"].items()}, "launches", d["gpu_launches"] // d["steps"], "cull", s["cull_tests"], s["cull_tests_by_family"], "packet", s["packet_tests"])
PY
tail -5 gpurun_out/exp1.err
