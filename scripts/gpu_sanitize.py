"""compute-sanitizer target: one small mixed scene (spheres, cubes, cylinders, a plane, reflections, two lights) and one CSG
scene, rendered once in each precision through the C ABI."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from go_raytracer_b200 import cuda, scenes
W, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (320, 200)
blobs = [scenes.to_blob(scenes.mixed(W, H, grid=(8, 4, 4), cones=True, depth=3)), scenes.to_blob(scenes.csg(W, H, groups=(3, 3)))]
for prec in (cuda.PRECISION_F32, cuda.PRECISION_F64_PARITY):
    ctx = cuda.Context((0,), prec)
    for blob in blobs:
        img, st = ctx.render(blob)
        print("precision", prec, "launches", st.kernel_launches, "rays", st.rays_primary, "checksum", int(img.astype("uint64").sum()), flush=True)
    ctx.close()
