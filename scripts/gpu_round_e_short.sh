#!/bin/bash
# r01e visit: parity tests, bench both arms, every BASELINE config once, ncu launch list of the bench command,
# (short form: without the ncu --set full capture)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench.json"))
print("value", d["value"], "ms", d["ms_per_step"], "serial", d["serialised_frame_ms"], "e2e", d["e2e"]["frame_ms"], d["e2e"]["phase_ms"], d["kernel_ms_per_frame"], d["cpu_baseline"], {k: d["roofline"].get(k) for k in ("kernel", "achieved", "peak", "frac", "traffic")})
PY
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
: > gpurun_out/configs.jsonl
for c in csg256 proctex mixed4096; do
  timeout 600 python bench.py --config $c --steps 5 --warmup 3 --cpu-crop 480 260 > gpurun_out/cfg_$c.out 2> gpurun_out/cfg_$c.err; echo "$c rc=$?"
  tail -2 gpurun_out/cfg_$c.err; tail -1 gpurun_out/cfg_$c.out >> gpurun_out/configs.jsonl
done
python - <<'PY'
import json
for l in open("gpurun_out/configs.jsonl"):
    try: d = json.loads(l)
    except Exception: continue
    print(d["config"]["workload"], "value", round(d["value"]), "ms", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["frame_ms"], 2), d["e2e"]["upload_ms_each"], {k: round(v, 2) for k, v in d["kernel_ms_per_frame"].items()}, d.get("cpu_baseline", {}).get("value"))
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/launches.csv \
   python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
