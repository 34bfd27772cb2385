#!/bin/bash
# multi-GPU visit (gpurun --gpus 8): config 2 at 2, 4, 8 ranks; config 5 (8K mixed4096) at 8 ranks (N = 1 comes from the 1-GPU visits)
mkdir -p gpurun_out; : > gpurun_out/scale.jsonl
tr() { n=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n --no-cpu-baseline "$@" >> gpurun_out/scale.jsonl 2> gpurun_out/scale_$n.err; echo "n=$n $* rc=$?"; }
for n in 2 4 8; do tr $n --steps 10 --warmup 3; done
tr 8 --config mixed4096 --steps 3 --warmup 3
python - <<'PY'
import json
for l in open("gpurun_out/scale.jsonl"):
    try: d = json.loads(l)
    except Exception: continue
    print(d["n_gpus"], d["config"]["workload"][:40], "value", round(d["value"]), "ms", round(d["ms_per_step"], 3), "e2e ms", round(d["e2e"]["frame_ms"], 3))
PY
for f in gpurun_out/scale_*.err; do grep -v "OMP_NUM_THREADS\|\*\*\*\*" "$f" | tail -n 2; done
