"""ctypes loader for the CPU oracle (oracle/oracle.c).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs - never by the go_raytracer_b200 package."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class Stats(ctypes.Structure):
    _fields_ = [("ms_total", ctypes.c_double), ("ms_kernels", ctypes.c_double), ("ms_h2d", ctypes.c_double),
                ("ms_d2h", ctypes.c_double), ("rays_primary", ctypes.c_uint64), ("rays_reflect", ctypes.c_uint64),
                ("rays_refract", ctypes.c_uint64), ("rays_shadow", ctypes.c_uint64), ("shadow_occluded", ctypes.c_uint64),
                ("leaf_tests", ctypes.c_uint64), ("flops_credited", ctypes.c_double), ("kernel_launches", ctypes.c_uint64),
                ("n_devices", ctypes.c_uint32), ("n_leaves", ctypes.c_uint32),
                ("cull_tests", ctypes.c_uint64), ("exact_tests", ctypes.c_uint64), ("flops_executed", ctypes.c_double),
                ("ms_class", ctypes.c_double * 9), ("launches_class", ctypes.c_uint64 * 9),
                ("packet_tests", ctypes.c_uint64), ("cull_tests_family", ctypes.c_uint64 * 3),
                ("rays_hit", ctypes.c_uint64), ("exact_tests_shadow", ctypes.c_uint64),
                ("rays_hit_primary", ctypes.c_uint64), ("list_tests", ctypes.c_uint64)]     # = gmlrt_stats, include/gmlrt.h

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}

    @property
    def rays_total(self):
        return self.rays_primary + self.rays_reflect + self.rays_refract + self.rays_shadow


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("oracle.c", "gomath.h")] + [os.path.join(_HERE, "..", "include", "gmlrt.h")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs if os.path.exists(s)):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = ctypes.CDLL(build())
        d, p = ctypes.c_double, ctypes.c_void_p
        L.oracle_render.restype = ctypes.c_int
        L.oracle_render.argtypes = [p, ctypes.c_size_t, p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                    ctypes.c_int, ctypes.c_int, ctypes.POINTER(Stats), p]
        L.oracle_render2.restype = ctypes.c_int
        L.oracle_render2.argtypes = L.oracle_render.argtypes + [p]
        L.oracle_ssim.restype = ctypes.c_int
        L.oracle_ssim.argtypes = [p, p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(d)]
        L.oracle_eval_surface.restype = ctypes.c_int
        L.oracle_eval_surface.argtypes = [p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_int, d, d, ctypes.POINTER(d)]
        L.oracle_intersect.restype = ctypes.c_int
        L.oracle_intersect.argtypes = [ctypes.c_int] + [ctypes.POINTER(d)] * 5 + [ctypes.POINTER(ctypes.c_int)]
        L.oracle_normal.restype = ctypes.c_int
        L.oracle_normal.argtypes = [ctypes.c_int, ctypes.POINTER(d), ctypes.POINTER(d), ctypes.c_int, ctypes.POINTER(d),
                                    ctypes.POINTER(d)]
        for name in ("tan", "sin", "cos", "acos"):
            f = getattr(L, "oracle_go_" + name)
            f.restype, f.argtypes = d, [d]
        for name in ("pow", "atan2"):
            f = getattr(L, "oracle_go_" + name)
            f.restype, f.argtypes = d, [d, d]
        L.oracle_pcg.restype = None
        L.oracle_pcg.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int, ctypes.POINTER(ctypes.c_uint64)]
        _LIB = L
    return _LIB


class OracleError(RuntimeError):
    def __init__(self, code):
        super().__init__(f"oracle error {code}")
        self.code = code


def render(blob, threads=8, crop=None, want_ids=False, want_chaos=False):
    """-> (rgba uint8 [H,W,4], Stats[, ids int32 [H,W,4]][, chaos bool [H,W]]).  crop = (x0, y0, x1, y1).
    chaos: pixels whose value depends on a cube face that rounding alone rejected (oracle.c cube case)."""
    import struct
    w, h = struct.unpack_from("<ii", blob, 16)
    out = np.zeros((h, w, 4), dtype=np.uint8)
    ids = np.full((h, w, 4), -2, dtype=np.int32) if want_ids else None
    st = Stats()
    x0, y0, x1, y1 = crop if crop else (0, 0, w, h)
    buf = ctypes.create_string_buffer(bytes(blob), len(blob))
    chaos = np.zeros((h, w), dtype=np.uint8) if want_chaos else None
    rc = lib().oracle_render2(buf, len(blob), out.ctypes.data, out.nbytes, threads, x0, y0, x1, y1, ctypes.byref(st),
                              ids.ctypes.data if want_ids else None, chaos.ctypes.data if want_chaos else None)
    if rc:
        raise OracleError(rc)
    res = (out, st) + ((ids,) if want_ids else ()) + ((chaos.astype(bool),) if want_chaos else ())
    return res


def eval_surface(blob, surface, face, u, v):
    m = (ctypes.c_double * 10)()
    buf = ctypes.create_string_buffer(bytes(blob), len(blob))
    rc = lib().oracle_eval_surface(buf, len(blob), surface, face, u, v, m)
    if rc:
        raise OracleError(rc)
    return tuple(m)


def ssim(a, b):
    """prim.SSIM (internal/prim/ssim.go:27-144) of two uint8 [H,W,4] rasters -> float."""
    a = np.ascontiguousarray(a, dtype=np.uint8)
    b = np.ascontiguousarray(b, dtype=np.uint8)
    if a.shape != b.shape:
        raise ValueError("images are not the same size")
    out = ctypes.c_double()
    rc = lib().oracle_ssim(a.ctypes.data, b.ctypes.data, a.shape[1], a.shape[0], ctypes.byref(out))
    if rc:
        raise ValueError("images are too small")
    return out.value


def stable_mask(blob, crop, threads=8, base=None):
    """Pixels of the window whose REFERENCE value survives a 1-ulp change of the field of view (-> bool [h, w]).
    Cube.Intersect bounds-checks a face hit on the face's own axis with no epsilon (raytracer.go:228), so a face is
    randomly missed when o + d t lands at -1e-17: rounding chaos in the reference itself (two float64 builds
    that differ by one ulp disagree there).  No float32 run can reproduce it, so scenes with cubes are compared on the
    stable pixels and the size of the mask is reported."""
    import struct
    x0, y0, x1, y1 = crop
    if base is None:
        base, _ = render(blob, threads=threads, crop=crop)
    b2 = bytearray(blob)
    fov = struct.unpack_from("<d", b2, 32)[0]
    struct.pack_into("<d", b2, 32, float(np.nextafter(fov, 1000.0)))
    pert, _ = render(bytes(b2), threads=threads, crop=crop)
    d = np.abs(base[y0:y1, x0:x1, :3].astype(np.int32) - pert[y0:y1, x0:x1, :3].astype(np.int32)).max(axis=2)
    return d <= 1


def parity_window(blob, img, ids, crop, threads=8, chaos=False):
    """The checker for throughput-mode frames at sizes the oracle cannot render whole: render the window
    `crop` = (x0, y0, x1, y1) of the same blob with the oracle (all 4 samples per pixel, the reference's own PCG
    streams) and compare it with the same window of a device frame `img` (uint8 [H,W,4]) and, when given, the device's
    per-sample primary hit ids `ids` (int32 [H,W,4], scan order, -1 = miss).  The north_star's bar: <= 1/255 per
    channel on >= 99.9% of pixels, zero hit/miss flips on primary rays (raytracer_test.go:37-52 is the reference's
    own, looser, acceptance).  -> (dict of plain numbers, oracle Stats, bad-pixel mask of the window)."""
    import time
    x0, y0, x1, y1 = crop
    t0 = time.perf_counter()
    oimg, ost, oids, ochaos = render(blob, threads=threads, crop=crop, want_ids=True, want_chaos=True)
    dt = time.perf_counter() - t0
    a = img[y0:y1, x0:x1, :3].astype(np.int32)
    b = oimg[y0:y1, x0:x1, :3].astype(np.int32)
    d = np.abs(a - b).max(axis=2)
    out = {"crop": [int(x0), int(y0), int(x1), int(y1)], "pixels": int(d.size), "within_1_255": float((d <= 1).mean()),
           "exact": float((d == 0).mean()), "max_abs_diff": int(d.max()), "oracle_seconds": round(dt, 2), "oracle_threads": int(threads)}
    if chaos:       # scenes with cubes: agreement where the reference's value does not hang on a face that rounding alone rejected
        st = ~ochaos[y0:y1, x0:x1]
        out["stable_fraction"] = float(st.mean())
        out["within_1_255_stable"] = float((d[st] <= 1).mean()) if st.any() else 0.0
    if ids is not None:
        gi, oi = ids[y0:y1, x0:x1], oids[y0:y1, x0:x1]
        out["primary_samples"] = int(gi.size)
        out["primary_flips"] = int(((gi < 0) != (oi < 0)).sum())
        out["id_mismatches"] = int((gi != oi).sum())
        if chaos:   # ... and the flips that are not the reference's own face-rejection chaos
            out["primary_flips_stable"] = int((((gi < 0) != (oi < 0)) & st[..., None]).sum())
            out["id_mismatches_stable"] = int(((gi != oi) & st[..., None]).sum())
    return out, ost, d > 1
