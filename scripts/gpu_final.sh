#!/bin/bash
# round-2 final evidence on one box: GPU tests, the default bench line (as the driver runs it), the reference arm, configs 3-5,
# parity mode, one frame under ncu --set full (after its plain run has exited 0), launch lists of a whole frame and of an eighth of it
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
T=${TAG:-r02z}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/${T}_smi.txt 2>&1; nproc >> gpurun_out/${T}_smi.txt
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/${T}_pytest.log | cut -c1-300
timeout 600 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_ref.json 2>> gpurun_out/${T}_bench.err; echo "ref rc=$?"
for c in csg256 proctex mixed4096; do
  timeout 600 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline >> gpurun_out/${T}_configs_3_4_5.jsonl 2>> gpurun_out/${T}_configs.err
done
timeout 600 python bench.py --precision f64 --steps 5 --warmup 3 --no-cpu-baseline --no-parity --no-other-configs > gpurun_out/${T}_bench_f64_parity_mode.json 2>> gpurun_out/${T}_bench.err
timeout 300 python scripts/gpu_one_frame.py spheres1024 1 && \
timeout 900 ncu --set full --clock-control none --profile-from-start off -o gpurun_out/${T}_frame_full -f python scripts/gpu_one_frame.py spheres1024 1 > gpurun_out/${T}_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/${T}_launches_spheres1024_4k.csv python scripts/gpu_one_frame.py spheres1024 1 > /dev/null 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/${T}_launches_one_eighth_share.csv python scripts/gpu_one_frame.py spheres1024 8 > /dev/null 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/${T}_launches_mixed4096_8k.csv python scripts/gpu_one_frame.py mixed4096 1 > /dev/null 2>&1
python - <<PY
import json
b=json.loads(open('gpurun_out/${T}_bench.json').read().strip().splitlines()[-1])
print({k:b.get(k) for k in ('value','ms_per_step','gpu_launches','clocks')}); print(b['e2e']['frame_ms'], b['e2e_new_scene']['frame_ms'], b['e2e_pageable']['frame_ms']); print({k:v for k,v in b['roofline'].items() if k not in ('note','by_class')}); print(b['parity']); print(b['cpu_baseline'])
print(open('gpurun_out/${T}_bench_ref.json').read()[:600])
for l in open('gpurun_out/${T}_configs_3_4_5.jsonl'):
    c=json.loads(l); print(c['config']['workload'], round(c['ms_per_step'],2), c.get('parity'))
f=json.loads(open('gpurun_out/${T}_bench_f64_parity_mode.json').read().strip().splitlines()[-1]); print('f64', f['ms_per_step'], f['value'])
PY
