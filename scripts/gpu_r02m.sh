#!/bin/bash
# round 2, state check after the container was re-created: GPU tests, bench (default line), configs 3-5, launch list
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r02m_smi.txt 2>&1
nproc >> gpurun_out/r02m_smi.txt
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02m_pytest.log 2>&1
tail -8 gpurun_out/r02m_pytest.log | cut -c1-300
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r02m_bench.json 2> gpurun_out/r02m_bench.err
tail -c 1500 gpurun_out/r02m_bench.err
for c in csg256 proctex mixed4096; do
  timeout 600 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline >> gpurun_out/r02m_configs_3_4_5.jsonl 2>> gpurun_out/r02m_configs.err
done
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-parity"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02m_launches.csv $CMD > gpurun_out/r02m_ncu_list.log 2>&1
python - <<'PY'
import json
b=json.loads(open('gpurun_out/r02m_bench.json').read().strip().splitlines()[-1])
print({k:b.get(k) for k in ('value','ms_per_step','e2e','roofline','parity','gpu_launches')})
print(b.get('kernel_ms_per_frame'))
for l in open('gpurun_out/r02m_configs_3_4_5.jsonl'):
    c=json.loads(l); print(c['config']['workload'], c['ms_per_step'], c.get('parity'))
PY
