#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/gpu_parts_balance.py 8 mixed4096 8 32 135 540 2>&1 | grep band
GMLRT_KEY_K=3 timeout 600 python scripts/gpu_parts_balance.py 8 mixed4096 8 135 2>&1 | grep band
GMLRT_KEY_K=2 timeout 600 python scripts/gpu_parts_balance.py 8 mixed4096 8 135 2>&1 | grep band
timeout 600 python scripts/gpu_parts_balance.py 8 spheres1024 8 32 135 270 2>&1 | grep band
GMLRT_KEY_K=3 timeout 600 python scripts/gpu_parts_balance.py 8 spheres1024 8 135 2>&1 | grep band
timeout 600 python scripts/gpu_parts_balance.py 2 spheres1024 8 135 540 1080 2>&1 | grep band
