#!/bin/bash
# experiment: big default tiles + level overflow retry (parity), CSG index-swap stack, all configs; ncu launch list at parts=8
mkdir -p gpurun_out; : > gpurun_out/exp3.jsonl; : > gpurun_out/exp3.err
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
run() { cfg=$1; shift; echo "== $cfg $*"; env "$@" timeout 300 python bench.py --config $cfg --steps 3 --warmup 3 --no-cpu-baseline 2>> gpurun_out/exp3.err | tail -1 >> gpurun_out/exp3.jsonl; }
run spheres1024 A=1
run mixed4096 A=1
run mixed4096 GMLRT_BRANCH_CAP=1.25
run csg256 A=1
run proctex A=1
python - <<'PY'
import json
for l in open("gpurun_out/exp3.jsonl"):
    try: d = json.loads(l)
    except Exception: print("bad line", l[:200]); continue
    s = d["roofline"]["scan_counters_per_frame"]
    print(d["config"]["workload"][:60], "ms", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["frame_ms"], 2), {k: round(v, 2) for k, v in d["kernel_ms_per_frame"].items()}, "launches", d["gpu_launches"] // d["steps"], "cull", s["cull_tests"], s["cull_tests_by_family"], "packet", s["packet_tests"])
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_parts8.csv \
   python bench.py --parts 8 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_parts8.log 2>&1; echo "ncu rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_csg.csv \
   python bench.py --config csg256 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_csg.log 2>&1; echo "ncu rc=$?"
tail -5 gpurun_out/exp3.err
