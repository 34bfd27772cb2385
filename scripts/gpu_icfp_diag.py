"""Developer diagnostic: ICFP fixture scenes, parity mode against the oracle, with the class-list debug switches."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from go_raytracer_b200 import cuda, lowering
from util import scene_blobs
import oracle
names = sys.argv[1:] or ["chess"]
for name in names:
    for file, blob in scene_blobs(name):
        want, ost, oids = oracle.render(blob, threads=os.cpu_count(), want_ids=True)
        for env in ({}, {"GMLRT_NO_CLASS_BOUNDS": "1"}, {"GMLRT_NO_SCREEN": "1"}):
            for k, v in env.items(): os.environ[k] = v
            ctx = cuda.Context((0,), cuda.PRECISION_F64_PARITY); ctx.set_counting(True)
            got, st = ctx.render(blob); ids = ctx.primary_hits(); ctx.close()
            for k in env: del os.environ[k]
            bad = (got != want).any(axis=2)
            idd = (ids != oids)
            print(name, file, env, "bad px", int(bad.sum()), "of", bad.size, "iddiff", int(idd.sum()), "rays", st.rays_total, ost.rays_total,
                  "hit", st.rays_hit, ost.rays_hit, "shadow", st.rays_shadow, ost.rays_shadow, "occl", st.shadow_occluded, ost.shadow_occluded, flush=True)
            ys, xs = np.nonzero(bad)
            for y, x in list(zip(ys, xs))[:8]:
                print("   ", y, x, "gpu", got[y, x, :3], "cpu", want[y, x, :3], "ids gpu", ids[y, x], "cpu", oids[y, x])
