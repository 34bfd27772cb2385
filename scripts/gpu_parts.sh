#!/bin/bash
# 1-GPU visit: parity tests, the bench line, and one rank's share of a 2/4/8-GPU frame (bench.py --parts)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -n 3 gpurun_out/bench.err
: > gpurun_out/parts.jsonl
for p in 2 4 8; do timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --parts $p >> gpurun_out/parts.jsonl 2>> gpurun_out/parts.err; done
python - <<'PY'
import json
for f in ("gpurun_out/bench.json", "gpurun_out/parts.jsonl"):
    for l in open(f):
        try: d = json.loads(l)
        except Exception: continue
        print(d["config"]["workload"][-28:], "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["frame_ms"], 2), d["e2e"]["phase_ms"], {k: round(v, 3) for k, v in d["kernel_ms_per_frame"].items()}, d["clocks"], "frac", round(d["roofline"]["frac"], 3))
PY
