"""Throughput-mode (float32) frames of BASELINE.json configs 2-5 at their stated sizes against the oracle on a window
(GPU box).  Prints one JSON line per config; writes the bad-pixel masks to gpurun_out/ for a look."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle
from go_raytracer_b200 import cuda, imageio, scenes

sys.path.insert(0, os.path.join(ROOT, "tests"))
from baseline_windows import CHAOS, WINDOWS, window_of

only = sys.argv[1:] or list(WINDOWS)
ctx = cuda.Context((0,), cuda.PRECISION_F32)
ctx.set_counting(True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
for name in only:
    gen, (W, H), frac, size = WINDOWS[name]
    t0 = time.time()
    blob = scenes.to_blob(gen(W, H))
    t1 = time.time()
    img, st = ctx.render(blob)
    ids = ctx.primary_hits()
    crop = window_of(W, H, frac, size)
    res, ost, bad = oracle.parity_window(blob, img, ids, crop, threads=os.cpu_count() or 8, chaos=name in CHAOS)
    res.update({"config": name, "size": [W, H], "lower_s": round(t1 - t0, 1), "gpu_ms": st.ms_kernels})
    print(json.dumps(res), flush=True)
    x0, y0, x1, y1 = crop
    vis = img[y0:y1, x0:x1].copy()
    vis[bad] = (255, 0, 255, 255)
    imageio.write_png(os.path.join(ROOT, "gpurun_out", f"parity_{name}.png"), vis)
    gi = ids[y0:y1, x0:x1]
ctx.close()
