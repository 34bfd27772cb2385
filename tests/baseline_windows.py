"""BASELINE.json configs 2-5 at their stated sizes and the window of each frame the oracle re-renders for the
float32 tolerance check (tests/test_gpu_parity.py, scripts/gpu_parity_full.py).  A window is given by its centre as a
fraction of the frame and its size in pixels; it is placed where the frame has silhouettes, shadows and texture edges
(thumbnails of the four frames were looked at), and sized so the oracle finishes it in well under a minute."""
from go_raytracer_b200 import scenes

WINDOWS = {
    #  name          generator                 frame size     window centre   window size
    "spheres1024": (scenes.spheres_lattice, (3840, 2160), (0.50, 0.50), (1280, 720)),
    "csg256":      (scenes.csg,             (3840, 2160), (0.50, 0.54), (1280, 720)),
    "proctex":     (scenes.proctex,         (3840, 2160), (0.50, 0.56), (1280, 720)),
    "mixed4096":   (scenes.mixed,           (7680, 4320), (0.50, 0.54), (640, 360)),
}
# scenes with cubes carry the reference's own rounding chaos (oracle.stable_mask): compared on the stable pixels
CHAOS = {"mixed4096"}


def window_of(W, H, centre, size):
    cw, ch = size
    x0 = max(0, min(W - cw, int(W * centre[0]) - cw // 2))
    y0 = max(0, min(H - ch, int(H * centre[1]) - ch // 2))
    return (x0, y0, x0 + cw, y0 + ch)
