"""ncu target: upload + 3 warm frames, then ONE frame between cudaProfilerStart / Stop (ncu --profile-from-start off).
usage: gpu_one_frame.py [config] [parts]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from go_raytracer_b200 import cuda, scenes
cfg = sys.argv[1] if len(sys.argv) > 1 else "spheres1024"
parts = int(sys.argv[2]) if len(sys.argv) > 2 else 1
text = {"spheres1024": scenes.spheres_lattice, "csg256": scenes.csg, "proctex": scenes.proctex, "mixed4096": scenes.mixed}[cfg]()
blob = scenes.to_blob(text)
ctx = cuda.Context((0,), cuda.PRECISION_F32)
if parts > 1:
    ctx.set_partition(0, parts)
ctx.upload(blob)
for _ in range(3):
    ctx.trace()
torch.cuda.synchronize()
rt = torch.cuda.cudart()
rt.cudaProfilerStart()
st = ctx.trace()
torch.cuda.synchronize()
rt.cudaProfilerStop()
print(cfg, "parts", parts, "launches", st.kernel_launches, "ms", st.ms_total, flush=True)
ctx.close()
