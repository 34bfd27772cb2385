"""Where does gmlrt_upload spend its time?  (GMLRT_DEBUG_UPLOAD=1 prints the phases.)"""
import os, sys, time
sys.path.insert(0, ".")
os.environ["GMLRT_DEBUG_UPLOAD"] = "1"
import numpy as np
from go_raytracer_b200 import cuda, scenes
blob = scenes.to_blob(scenes.spheres_lattice(3840, 2160, depth=3))
c = cuda.Context((0,), cuda.PRECISION_F32)
out = np.empty((2160, 3840, 4), np.uint8)
for k in range(8):
    t0 = time.perf_counter(); c.upload(blob); t1 = time.perf_counter(); c.trace(); t2 = time.perf_counter(); c.download(out); t3 = time.perf_counter()
    print(f"iter {k}: upload {1e3*(t1-t0):.3f} trace {1e3*(t2-t1):.3f} download {1e3*(t3-t2):.3f} ms", file=sys.stderr)
