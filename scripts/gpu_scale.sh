#!/bin/bash
# multi-GPU visit (gpurun --gpus N): parity tests on GPU 0, then bench.py at 1, 2, .. N ranks (both gather modes at N)
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
: > gpurun_out/scale.jsonl
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline >> gpurun_out/scale.jsonl 2> gpurun_out/scale_1.err; echo "n=1 rc=$?"
for n in 2 4 8; do
  [ $n -le $N ] || continue
  for g in ipc nccl; do
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --steps 10 --warmup 3 --no-cpu-baseline --gather $g >> gpurun_out/scale.jsonl 2> gpurun_out/scale_${n}_$g.err; echo "n=$n $g rc=$?"
  done
done
# in-process multi-device context (gmlrt_create with N devices): parity vs 1 device + timing
timeout 300 python scripts/gpu_multidev.py $N > gpurun_out/multidev.log 2>&1; echo "multidev rc=$?"; tail -5 gpurun_out/multidev.log
python - <<'PY'
import json
for l in open("gpurun_out/scale.jsonl"):
    try: d = json.loads(l)
    except Exception: continue
    print(d["n_gpus"], d["config"]["parallelism"][:60], "value", round(d["value"]), "ms", round(d["ms_per_step"], 3), "e2e ms", round(d["e2e"]["frame_ms"], 3))
PY
for f in gpurun_out/scale_*.err; do tail -n 3 "$f"; done
