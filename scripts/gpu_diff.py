"""Developer diagnostic: where does float32 mode differ from the oracle?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from go_raytracer_b200 import cuda, scenes, imageio, lowering
from util import scene_blobs
import oracle
os.makedirs("gpurun_out", exist_ok=True)
cases = {"mixed": scenes.to_blob(scenes.mixed(320, 180, grid=(8, 8, 4), depth=4)),
         "proctex": scenes.to_blob(scenes.proctex(320, 180, n_spheres=16)),
         "cube": scene_blobs("cube")[0][1]}
ctx = cuda.Context((0,), cuda.PRECISION_F32); ctx.set_counting(True)
for name, blob in cases.items():
    oimg, ost, oids = oracle.render(blob, threads=os.cpu_count(), want_ids=True)
    img, st = ctx.render(blob)
    ids = ctx.primary_hits()
    d = np.abs(img[..., :3].astype(int) - oimg[..., :3].astype(int)).max(axis=2)
    bad = d > 1
    flips = ((ids < 0) != (oids < 0))
    iddiff = (ids != oids)
    d_nodes = lowering.describe(blob)
    print(name, "bad px", int(bad.sum()), "of", bad.size, "flips", int(flips.sum()), "iddiff samples", int(iddiff.sum()),
          "bad px with iddiff", int((bad & iddiff.any(axis=2)).sum()), "hist", np.bincount(np.minimum(d[bad], 50) // 10).tolist())
    # kinds of the objects involved in bad pixels (oracle's primary id of sample 0)
    leaves = [n for n in d_nodes["nodes"] if n["kind"] >= 16]
    ys, xs = np.nonzero(bad)
    for y, x in list(zip(ys, xs))[:12]:
        print("   ", y, x, "gpu", img[y, x, :3], "cpu", oimg[y, x, :3], "ids gpu", ids[y, x], "cpu", oids[y, x])
    vis = np.zeros_like(oimg); vis[..., 3] = 255; vis[bad] = (255, 0, 0, 255)
    imageio.write_png(f"gpurun_out/diff_{name}.png", np.concatenate([img, oimg, vis], axis=1))
