"""Developer diagnostic: float32 agreement / flips on the ICFP fixture scenes and the cube golden (what the relaxed bars of the tests hold)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from go_raytracer_b200 import cuda
from util import scene_blobs
import oracle
names = ["checked-cube", "chess", "cone", "cone-fractal", "cube2", "dice", "ellipsoid", "fov", "fractal", "golf", "holes", "house",
         "intercyl", "large", "pipe", "rotate", "snowgoon", "spheres", "spotlight", "cube"]
ctx = cuda.Context((0,), cuda.PRECISION_F32); ctx.set_counting(True)
worst = 1.0
for name in names:
    for file, blob in scene_blobs(name):
        want, ost, oids, chaos = oracle.render(blob, threads=os.cpu_count(), want_ids=True, want_chaos=True)
        got, _ = ctx.render(blob); ids = ctx.primary_hits()
        d = np.abs(got[..., :3].astype(int) - want[..., :3].astype(int)).max(axis=2)
        st = ~chaos
        w = float((d[st] <= 1).mean()); fl = int((((ids < 0) != (oids < 0)) & st[..., None]).sum())
        worst = min(worst, w)
        print(name, file, "stable", round(float(st.mean()), 4), "within(stable)", round(w, 5), "flips(stable)", fl, flush=True)
print("worst", worst)
