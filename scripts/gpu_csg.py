"""Developer check on a GPU box: CSG scene (extension) in both modes vs the CPU oracle."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from go_raytracer_b200 import cuda, scenes, imageio
from util import agreement
import oracle
for name, text in (("csg", scenes.csg(640, 360, groups=(8, 8))), ("mixed", scenes.mixed(640, 360, grid=(8, 8, 4), depth=3))):
    blob = scenes.to_blob(text)
    oimg, ost, oids = oracle.render(blob, threads=os.cpu_count(), want_ids=True)
    for prec, pname in ((cuda.PRECISION_F64_PARITY, "f64"), (cuda.PRECISION_F32, "f32")):
        ctx = cuda.Context((0,), prec)
        ctx.set_counting(True)
        t = time.time(); img, st = ctx.render(blob); dt = time.time() - t
        ids = ctx.primary_hits()
        print(name, pname, f"{dt*1e3:.1f} ms kernels {st.ms_kernels:.2f} exact {float((img==oimg).all(axis=2).mean()):.6f} within1 {agreement(img, oimg):.6f}",
              "flips", int(((ids < 0) != (oids < 0)).sum()), "iddiff", int((ids != oids).sum()),
              "rays", st.rays_total, ost.rays_total, "flops", st.flops_credited, ost.flops_credited, "tests", st.leaf_tests, ost.leaf_tests,
              "cull", st.cull_tests, "exact", st.exact_tests, flush=True)
        ctx.close()
        os.makedirs("gpurun_out", exist_ok=True)
        imageio.write_png(f"gpurun_out/{name}_{pname}.png", img)
