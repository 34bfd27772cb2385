#!/bin/bash
# round 2, first trip: where does float32 stand at the benched sizes; baseline numbers for this round
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r02a_smi.txt 2>&1
nproc >> gpurun_out/r02a_smi.txt
timeout 1500 python scripts/gpu_parity_full.py > gpurun_out/r02a_parity.jsonl 2> gpurun_out/r02a_parity.err
timeout 900 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_parity.py::test_f32_tolerance_at_baseline_size > gpurun_out/r02a_pytest.log 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err
timeout 600 python bench.py --steps 5 --warmup 3 --precision f64 --no-cpu-baseline > gpurun_out/r02a_bench_f64.json 2> gpurun_out/r02a_bench_f64.err
tail -3 gpurun_out/r02a_pytest.log
cat gpurun_out/r02a_parity.jsonl
