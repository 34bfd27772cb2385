"""Where does gmlrt_upload spend its time?  (GMLRT_DEBUG_UPLOAD=1 prints the phases.)  usage: gpu_upload.py [config]"""
import os, sys, time
sys.path.insert(0, ".")
os.environ["GMLRT_DEBUG_UPLOAD"] = "1"
import numpy as np
from go_raytracer_b200 import cuda, scenes
config = sys.argv[1] if len(sys.argv) > 1 else "spheres1024"
W, H = (7680, 4320) if config == "mixed4096" else (3840, 2160)
text = {"spheres1024": lambda: scenes.spheres_lattice(W, H, depth=3), "csg256": lambda: scenes.csg(W, H, depth=2),
        "mixed4096": lambda: scenes.mixed(W, H, depth=5)}[config]()
blob = scenes.to_blob(text)
c = cuda.Context((0,), cuda.PRECISION_F32)
out = np.empty((H, W, 4), np.uint8)
for k in range(8):
    t0 = time.perf_counter(); c.upload(blob); t1 = time.perf_counter(); c.trace(); t2 = time.perf_counter(); c.download(out); t3 = time.perf_counter()
    print(f"{config} iter {k}: upload {1e3*(t1-t0):.3f} trace {1e3*(t2-t1):.3f} download {1e3*(t3-t2):.3f} ms", file=sys.stderr)
