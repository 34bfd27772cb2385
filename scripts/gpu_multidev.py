"""gmlrt_create on N devices of one process: rows dealt in 8-row bands, devices > 0 store into device 0's frame over
NVLink from the level-0 k_resolve.  Checks the raster against 1 device (parity mode: byte-exact, ids too) and times the f32 frame."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from go_raytracer_b200 import cuda, scenes

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
blob = scenes.to_blob(scenes.mixed(640, 363, grid=(6, 6, 4), depth=3))
one = cuda.Context((0,), cuda.PRECISION_F64_PARITY)
one.set_counting(True)
ref, rst = one.render(blob)
ref_ids = one.primary_hits()
many = cuda.Context(tuple(range(n)), cuda.PRECISION_F64_PARITY)
many.set_counting(True)
img, st = many.render(blob)
ids = many.primary_hits()
print("devices", st.n_devices, "raster equal", np.array_equal(img, ref), "ids equal", np.array_equal(ids, ref_ids),
      "rays equal", (st.rays_primary, st.rays_reflect, st.rays_refract, st.rays_shadow) == (rst.rays_primary, rst.rays_reflect, rst.rays_refract, rst.rays_shadow))
assert np.array_equal(img, ref) and np.array_equal(ids, ref_ids)
one.close(); many.close()

blob = scenes.to_blob(scenes.spheres_lattice(3840, 2160, depth=3))
for devs in [(0,), tuple(range(n))]:
    c = cuda.Context(devs, cuda.PRECISION_F32)
    c.upload(blob)
    for _ in range(3):
        c.trace()
    t0 = time.perf_counter()
    ms = [c.trace().ms_total for _ in range(10)]
    wall = (time.perf_counter() - t0) / 10 * 1e3
    print(f"{len(devs)} device(s): frame {np.median(ms):.3f} ms (device events), {wall:.3f} ms wall")
    c.close()
