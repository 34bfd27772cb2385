"""Where do upload and download spend their time?  (GPU box)"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from go_raytracer_b200 import cuda, scenes
blob = scenes.to_blob(scenes.spheres_lattice(3840, 2160))
ctx = cuda.Context((0,), cuda.PRECISION_F32)
for k in range(4):
    t = time.perf_counter(); ctx.upload(blob); print("upload %.3f ms" % ((time.perf_counter() - t) * 1e3), flush=True)
ctx.trace()
hf = cuda.HostFrame(3840, 2160)
tp = torch.empty((2160, 3840, 4), dtype=torch.uint8, pin_memory=True).numpy()
pg = np.empty((2160, 3840, 4), dtype=np.uint8); pg[...] = 0
for name, dst in (("HostFrame", hf.array), ("torch pinned", tp), ("pageable", pg)):
    for k in range(4):
        t = time.perf_counter(); ctx.download(dst); dt = (time.perf_counter() - t) * 1e3
    print("download into %s: %.3f ms" % (name, dt), flush=True)
