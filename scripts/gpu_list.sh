#!/bin/bash
# ncu launch list only (per-kernel durations), args: extra bench.py flags
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline $*"
timeout 280 $CMD > gpurun_out/plain1.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 230 -c 260 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
python - <<'PY'
import csv, collections, re
rows = list(csv.reader(open('gpurun_out/launches.csv')))
for i, r in enumerate(rows):
    if 'Kernel Name' in r: hdr = r; start = i + 1; break
ki, vi = hdr.index('Kernel Name'), hdr.index('Metric Value')
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[start:]:
    if len(r) <= vi: continue
    name = re.sub(r'\(.*', '', r[ki]).replace('gmlrt::', '')[:70]
    try: v = float(r[vi].replace(',', ''))
    except ValueError: continue
    agg[name][0] += 1; agg[name][1] += v
tot = sum(v[1] for v in agg.values())
print('total ns', tot)
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]): print(f"{v[1]/tot*100:5.1f}% {v[1]/1e3:10.1f} us n={v[0]:3d} {k}")
PY
