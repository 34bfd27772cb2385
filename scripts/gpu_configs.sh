#!/bin/bash
# bench every BASELINE.json config once (short), raw JSON lines into gpurun_out/configs.jsonl
mkdir -p gpurun_out; : > gpurun_out/configs.jsonl
for c in "$@"; do
  echo "== $c"
  timeout 900 python bench.py --config $c --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/cfg_$c.out 2> gpurun_out/cfg_$c.err
  echo "rc=$?"; tail -3 gpurun_out/cfg_$c.err; tail -1 gpurun_out/cfg_$c.out >> gpurun_out/configs.jsonl
done
