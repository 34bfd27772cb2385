#!/bin/bash
# experiment: region keys (parity + all configs), tile size on config 2, band height balance at 8 parts
mkdir -p gpurun_out; : > gpurun_out/exp2.jsonl; : > gpurun_out/exp2.err
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
run() { cfg=$1; shift; echo "== $cfg $*"; env "$@" timeout 300 python bench.py --config $cfg --steps 3 --warmup 3 --no-cpu-baseline 2>> gpurun_out/exp2.err | tail -1 >> gpurun_out/exp2.jsonl; }
run spheres1024 A=1
run spheres1024 GMLRT_POOL_MB=49152 GMLRT_TILE_SAMPLES=34000000
run mixed4096 GMLRT_POOL_MB=98304 GMLRT_TILE_SAMPLES=67108864
run mixed4096 GMLRT_POOL_MB=98304 GMLRT_TILE_SAMPLES=67108864 GMLRT_LARGE_CELLS=0.75
run mixed4096 A=1
run csg256 A=1
run proctex A=1
python - <<'PY'
import json
for l in open("gpurun_out/exp2.jsonl"):
    try: d = json.loads(l)
    except Exception: print("bad line", l[:200]); continue
    s = d["roofline"]["scan_counters_per_frame"]
    print(d["config"]["workload"][:60], "ms", round(d["ms_per_step"], 2), {k: round(v, 2) for k, v in d["kernel_ms_per_frame"].items()}, "launches", d["gpu_launches"] // d["steps"], "cull", s["cull_tests"], s["cull_tests_by_family"], "packet", s["packet_tests"])
PY
timeout 600 python scripts/gpu_parts_balance.py 8 2>&1 | tail -8
tail -5 gpurun_out/exp2.err
