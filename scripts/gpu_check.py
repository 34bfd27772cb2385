"""Developer smoke on a GPU box: render the golden scenes in both precision modes and report
agreement with the goldens and with the CPU oracle.  (Uses oracle/ => a test script.)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from go_raytracer_b200 import cuda
from util import GOLDENS, scene_blobs, golden, agreement
import oracle

names = sys.argv[1:] or ["cylinder", "sphere", "cube", "canned"]
for prec, pname in ((cuda.PRECISION_F64_PARITY, "f64"), (cuda.PRECISION_F32, "f32")):
    ctx = cuda.Context((0,), prec)
    ctx.set_counting(True)
    for name in names:
        for file, blob in scene_blobs(name):
            t = time.time(); img, st = ctx.render(blob); dt = time.time() - t
            want = golden(GOLDENS[name][file])
            exact = float((np.abs(img[..., :3].astype(int) - want[..., :3].astype(int)).max(axis=2) == 0).mean())
            oimg, ost, oids = oracle.render(blob, threads=os.cpu_count(), want_ids=True)
            ids = ctx.primary_hits()
            flips = int(((ids < 0) != (oids < 0)).sum())
            iddiff = int((ids != oids).sum())
            print(f"{pname} {file:14s} {dt*1e3:8.1f} ms  kernels {st.ms_kernels:8.2f} ms  exact {exact:.6f}  within1 {agreement(img, want):.6f} "
                  f"vs-oracle-exact {float((img==oimg).all(axis=2).mean()):.6f} flips {flips} iddiff {iddiff} "
                  f"rays gpu {st.rays_primary}/{st.rays_reflect}/{st.rays_refract}/{st.rays_shadow}/{st.shadow_occluded} "
                  f"cpu {ost.rays_primary}/{ost.rays_reflect}/{ost.rays_refract}/{ost.rays_shadow}/{ost.shadow_occluded} "
                  f"flops gpu {st.flops_credited:.4g} cpu {ost.flops_credited:.4g} launches {st.kernel_launches}", flush=True)
    ctx.close()
