#!/bin/bash
# usage: gpu_scale_n.sh N [extra bench args]: bench.py under torchrun on N GPUs of one box, as the driver launches it
cd "$GRAFT_REPO_ROOT" || exit 1
N=$1; TAG=${2:-cfg2}; shift; shift
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/scale_topo_$N.txt 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps ${STEPS:-20} --warmup 3 --verify "$@" > gpurun_out/scale_n${N}_$TAG.json 2> gpurun_out/scale_n${N}_$TAG.err
echo "rc=$?"; tail -c 600 gpurun_out/scale_n${N}_$TAG.err
python - <<PY
import json
b=json.loads(open('gpurun_out/scale_n${N}_$TAG.json').read().strip().splitlines()[-1])
print({k:b.get(k) for k in ('n_gpus','value','ms_per_step','frame_ms_with_barrier','gather_check')})
for k in ('e2e','e2e_ipc_gather','e2e_host_gather','e2e_new_scene','e2e_pageable'):
    if k in b: print(k, {kk:vv for kk,vv in b[k].items() if kk in ('frame_ms','phase_ms','gather','gather_check','destination')})
PY
