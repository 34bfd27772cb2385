"""Developer diagnostic: one ICFP scene in one precision against the oracle: id pairs, difference image."""
import os, sys, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from go_raytracer_b200 import cuda, lowering, imageio
from util import scene_blobs
import oracle
os.makedirs("gpurun_out", exist_ok=True)
for arg in sys.argv[1:]:
    name, prec = arg.split(":")
    for file, blob in scene_blobs(name):
        want, ost, oids = oracle.render(blob, threads=os.cpu_count(), want_ids=True)
        ctx = cuda.Context((0,), cuda.PRECISION_F64_PARITY if prec == "f64" else cuda.PRECISION_F32); ctx.set_counting(True)
        got, st = ctx.render(blob); ids = ctx.primary_hits(); ctx.close()
        d = np.abs(got[..., :3].astype(int) - want[..., :3].astype(int)).max(axis=2)
        bad = d > (0 if prec == "f64" else 1)
        idd = ids != oids
        print(name, file, prec, "bad px", int(bad.sum()), "of", bad.size, "iddiff", int(idd.sum()), "rays", st.rays_total, ost.rays_total,
              "hit", st.rays_hit, ost.rays_hit, "prim-hit", st.rays_hit_primary, int((oids >= 0).sum()), "refl", st.rays_reflect, ost.rays_reflect,
              "refr", st.rays_refract, ost.rays_refract, "shadow", st.rays_shadow, ost.rays_shadow, "occl", st.shadow_occluded, ost.shadow_occluded, flush=True)
        pairs = collections.Counter(zip(ids[idd].tolist(), oids[idd].tolist()))
        print("   id pairs (gpu, cpu): count", pairs.most_common(24))
        nodes = lowering.describe(blob)["nodes"]
        print("   nodes", len(nodes), "kinds", collections.Counter(n["kind"] for n in nodes))
        vis = np.zeros_like(want); vis[..., 3] = 255; vis[bad] = (255, 0, 0, 255); vis[idd.any(axis=2)] = (255, 255, 0, 255)
        imageio.write_png(f"gpurun_out/diag_{name}_{prec}.png", np.concatenate([got, want, vis], axis=1))
