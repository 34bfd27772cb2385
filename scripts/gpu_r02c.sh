#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_parity.py::test_f32_tolerance_at_baseline_size > gpurun_out/r02c_pytest.log 2>&1
tail -5 gpurun_out/r02c_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02c_bench.json 2> gpurun_out/r02c_bench.err
for c in csg256 proctex mixed4096; do timeout 600 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline >> gpurun_out/r02c_configs.jsonl 2>> gpurun_out/r02c_configs.err; done
python - <<'PY'
import json
for f in ("r02c_bench.json",):
    try:
        b=json.load(open("gpurun_out/"+f)); print(f, round(b["ms_per_step"],3), round(b["e2e"]["frame_ms"],3), {k:round(v,3) for k,v in b["kernel_ms_per_frame"].items()}, b["roofline"]["scan_counters_per_frame"])
    except Exception as e: print(f, "failed", e)
for l in open("gpurun_out/r02c_configs.jsonl"):
    b=json.loads(l); print(b["config"]["workload"], round(b["ms_per_step"],3), {k:round(v,3) for k,v in b["kernel_ms_per_frame"].items()})
PY
timeout 1500 python scripts/gpu_parity_full.py > gpurun_out/r02c_parity.jsonl 2> gpurun_out/r02c_parity.err
cat gpurun_out/r02c_parity.jsonl
GMLRT_PRIMARY_POINTS=0 timeout 600 python scripts/gpu_parity_full.py spheres1024 proctex > gpurun_out/r02c_parity_nopoints.jsonl 2>> gpurun_out/r02c_parity.err
cat gpurun_out/r02c_parity_nopoints.jsonl
