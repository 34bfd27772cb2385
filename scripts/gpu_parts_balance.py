"""One GPU stands in for each rank of an N-GPU frame in turn: per-part frame time for several band heights
(GMLRT_BAND_ROWS), to pick the band that minimises the SLOWEST part (= the N-GPU frame time).
usage: gpu_parts_balance.py N [config] [band ...]"""
import os, sys
sys.path.insert(0, ".")
import numpy as np
from go_raytracer_b200 import cuda, scenes

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
config = sys.argv[2] if len(sys.argv) > 2 else "spheres1024"
bands = [int(b) for b in sys.argv[3:]] or [8, 16, 32, 64, 135]
text = scenes.mixed(7680, 4320, depth=5) if config == "mixed4096" else scenes.spheres_lattice(3840, 2160, depth=3)
blob = scenes.to_blob(text)
c = cuda.Context((0,), cuda.PRECISION_F32)
for band in bands:
    os.environ["GMLRT_BAND_ROWS"] = str(band)
    ms = []
    for part in range(n):
        c.set_partition(part, n)
        c.upload(blob)
        for _ in range(2):
            c.trace()
        ms.append(float(np.median([c.trace().ms_total for _ in range(3)])))
    print(f"{config} K={os.environ.get('GMLRT_KEY_K', 'dflt')} band {band:4d}: max {max(ms):.3f} mean {np.mean(ms):.3f} ms  parts {[round(x, 2) for x in ms]}", flush=True)
c.close()
