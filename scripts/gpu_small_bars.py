"""Developer diagnostic: float32 agreement and primary flips on the small synthetic scenes of the test-suite."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from go_raytracer_b200 import cuda, scenes
import oracle
gens = {"spheres": scenes.spheres_lattice(320, 180, grid=(8, 4, 4)), "mixed": scenes.mixed(320, 180, grid=(8, 8, 4), depth=4),
        "proctex": scenes.proctex(320, 180, n_spheres=16), "csg": scenes.csg(320, 180, groups=(4, 4)),
        "smoke": scenes.mixed(320, 200, grid=(8, 4, 4), cones=False, depth=3)}
ctx = cuda.Context((0,), cuda.PRECISION_F32); ctx.set_counting(True)
for name, text in gens.items():
    blob = scenes.to_blob(text)
    want, ost, oids, chaos = oracle.render(blob, threads=os.cpu_count(), want_ids=True, want_chaos=True)
    got, st = ctx.render(blob); ids = ctx.primary_hits()
    d = np.abs(got[..., :3].astype(int) - want[..., :3].astype(int)).max(axis=2)
    fl = ((ids < 0) != (oids < 0))
    st_ = ~chaos
    print(name, "within", float((d <= 1).mean()), "stable frac", float(st_.mean()), "within stable", float((d[st_] <= 1).mean()), "flips", int(fl.sum()), "flips stable", int((fl & st_[..., None]).sum()), flush=True)
