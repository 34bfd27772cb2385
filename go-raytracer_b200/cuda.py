"""ctypes bridge to libgmlrt.so - the Python twin of the Go ``internal/cuda`` bridge the
north_star describes (cgo -> extern "C" -> CUDA runtime).  It owns nothing but a context
handle; device memory, streams and the copy-back live behind the C ABI (include/gmlrt.h).

There is NO CPU fallback: a missing library or a box without a CUDA device raises."""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GMLRT_LIB") or os.path.join(_HERE, "libgmlrt.so")      # GMLRT_LIB: another build of the same library (A/B measurements)

IPC_HANDLE_BYTES = 64
PRECISION_F32 = 0
PRECISION_F64_PARITY = 1

ERROR_NAMES = {0: "GMLRT_OK", -1: "GMLRT_E_BAD_ARG", -2: "GMLRT_E_BAD_BLOB", -3: "GMLRT_E_UNSUPPORTED", -4: "GMLRT_E_NO_DEVICE",
               -5: "GMLRT_E_CUDA", -6: "GMLRT_E_OOM", -7: "GMLRT_E_SURFACE"}


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
                ("rays_hit_primary", ctypes.c_uint64), ("list_tests", ctypes.c_uint64)]

    KERNEL_CLASSES = ("gen", "cull_primary", "cull_secondary", "cull_shadow", "shade", "resolve", "primary_points", "exact", "sort")

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}

    @property
    def rays_total(self):
        return self.rays_primary + self.rays_reflect + self.rays_refract + self.rays_shadow


class GmlrtError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"{ERROR_NAMES.get(code, code)}: {text}")
        self.code = code


EXPORTS = ("gmlrt_download_rows", "gmlrt_ssim_device_noalloc", "gmlrt_alloc_host", "gmlrt_free_host", "gmlrt_ssim", "gmlrt_frame_ipc_export", "gmlrt_frame_ipc_attach", "gmlrt_measure_fp32_tflops", "gmlrt_set_profiling", "gmlrt_set_partition", "gmlrt_frame_device", "gmlrt_create", "gmlrt_render", "gmlrt_upload", "gmlrt_trace", "gmlrt_download", "gmlrt_set_counting",
           "gmlrt_primary_hits", "gmlrt_stream", "gmlrt_last_error", "gmlrt_destroy", "gmlrt_blob_version", "gmlrt_build_info")

_lib = None


def load():
    """Load libgmlrt.so (built in-tree by __graft_entry__.build()).  Fails loudly if absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GmlrtError(-4, f"{LIB_PATH} is missing - run `python -c 'import __graft_entry__ as g; g.build()'`; "
                                 "there is no CPU fallback")
        L = ctypes.CDLL(LIB_PATH)
        p, sz, i = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int
        L.gmlrt_create.restype, L.gmlrt_create.argtypes = i, [ctypes.POINTER(p), ctypes.POINTER(i), i, ctypes.c_uint]
        L.gmlrt_render.restype, L.gmlrt_render.argtypes = i, [p, p, sz, p, sz, ctypes.POINTER(Stats)]
        L.gmlrt_upload.restype, L.gmlrt_upload.argtypes = i, [p, p, sz]
        L.gmlrt_trace.restype, L.gmlrt_trace.argtypes = i, [p, ctypes.POINTER(Stats)]
        L.gmlrt_download.restype, L.gmlrt_download.argtypes = i, [p, p, sz]
        L.gmlrt_download_rows.restype, L.gmlrt_download_rows.argtypes = i, [p, p, sz]
        L.gmlrt_set_counting.restype, L.gmlrt_set_counting.argtypes = i, [p, i]
        L.gmlrt_measure_fp32_tflops.restype, L.gmlrt_measure_fp32_tflops.argtypes = ctypes.c_double, [i]
        L.gmlrt_set_profiling.restype, L.gmlrt_set_profiling.argtypes = i, [p, i]
        L.gmlrt_primary_hits.restype, L.gmlrt_primary_hits.argtypes = i, [p, p, sz]
        L.gmlrt_stream.restype, L.gmlrt_stream.argtypes = p, [p]
        L.gmlrt_set_partition.restype, L.gmlrt_set_partition.argtypes = i, [p, i, i]
        L.gmlrt_frame_device.restype, L.gmlrt_frame_device.argtypes = p, [p]
        L.gmlrt_ssim.restype, L.gmlrt_ssim.argtypes = i, [p, p, p, i, i, ctypes.POINTER(ctypes.c_double)]
        L.gmlrt_ssim_device_noalloc.restype, L.gmlrt_ssim_device_noalloc.argtypes = i, [p, p, p, i, i, ctypes.POINTER(ctypes.c_double)]
        L.gmlrt_frame_ipc_export.restype, L.gmlrt_frame_ipc_export.argtypes = i, [p, p, sz]
        L.gmlrt_frame_ipc_attach.restype, L.gmlrt_frame_ipc_attach.argtypes = i, [p, p, sz]
        L.gmlrt_alloc_host.restype, L.gmlrt_alloc_host.argtypes = p, [sz]
        L.gmlrt_free_host.restype, L.gmlrt_free_host.argtypes = None, [p]
        L.gmlrt_last_error.restype, L.gmlrt_last_error.argtypes = ctypes.c_char_p, [p]
        L.gmlrt_destroy.restype, L.gmlrt_destroy.argtypes = None, [p]
        L.gmlrt_blob_version.restype, L.gmlrt_blob_version.argtypes = i, []
        L.gmlrt_build_info.restype, L.gmlrt_build_info.argtypes = ctypes.c_char_p, []
        _lib = L
    return _lib


def measure_fp32_tflops(device=0):
    """FFMA-chain microbenchmark on `device` (TFLOP/s): the measured FP32 roofline denominator."""
    return float(load().gmlrt_measure_fp32_tflops(int(device)))


class HostFrame:
    """A page-locked RGBA8 raster (gmlrt_alloc_host) viewed as a numpy array: the destination the bridge hands to
    gmlrt_render / gmlrt_download so that the copy-back is one direct DMA."""

    def __init__(self, width, height):
        L = load()
        self._L, self.nbytes = L, width * height * 4
        self._p = L.gmlrt_alloc_host(self.nbytes)
        if not self._p:
            raise GmlrtError(-6, "gmlrt_alloc_host failed")
        self.array = np.ctypeslib.as_array((ctypes.c_uint8 * self.nbytes).from_address(self._p)).reshape(height, width, 4)

    def free(self):
        if self._p:
            self.array = None
            self._L.gmlrt_free_host(self._p)
            self._p = None

    __del__ = free


class Context:
    """One gmlrt_ctx: a set of devices + a precision mode; reusable across renders."""

    def __init__(self, devices=(0,), precision=PRECISION_F32):
        L = load()
        self._L = L
        self._h = ctypes.c_void_p()
        devs = (ctypes.c_int * len(devices))(*devices)
        rc = L.gmlrt_create(ctypes.byref(self._h), devs, len(devices), precision)
        if rc:
            raise GmlrtError(rc, L.gmlrt_last_error(None).decode())
        self.devices, self.precision = tuple(devices), precision
        self._size = None

    def _check(self, rc):
        if rc:
            raise GmlrtError(rc, self._L.gmlrt_last_error(self._h).decode())

    def close(self):
        if self._h:
            self._L.gmlrt_destroy(self._h)
            self._h = None               # (not ctypes.c_void_p(): at interpreter shutdown the module may already be gone)

    __del__ = close

    @staticmethod
    def _blob_size(blob):
        import struct
        return struct.unpack_from("<ii", blob, 16)

    def render(self, blob, out=None):
        """Host blob in, host raster out (H2D + kernels + D2H).  -> (uint8 [H,W,4], Stats)."""
        w, h = self._blob_size(blob)
        if out is None:
            out = np.empty((h, w, 4), dtype=np.uint8)
        st = Stats()
        buf = (ctypes.c_char * len(blob)).from_buffer_copy(blob) if not isinstance(blob, ctypes.Array) else blob
        self._check(self._L.gmlrt_render(self._h, buf, len(blob), out.ctypes.data, out.nbytes, ctypes.byref(st)))
        self._size = (w, h)
        return out, st

    def upload(self, blob):
        buf = (ctypes.c_char * len(blob)).from_buffer_copy(blob)
        self._check(self._L.gmlrt_upload(self._h, buf, len(blob)))
        self._size = self._blob_size(blob)

    def upload_from(self, address, nbytes, size):
        """gmlrt_upload of a blob that already sits at a host address (e.g. page-locked memory from gmlrt_alloc_host, so that
        the copy to the device is one DMA); size = (width, height) of its frame."""
        self._check(self._L.gmlrt_upload(self._h, ctypes.c_void_p(address), nbytes))
        self._size = tuple(size)

    def trace(self):
        st = Stats()
        self._check(self._L.gmlrt_trace(self._h, ctypes.byref(st)))
        return st

    def download(self, out=None):
        w, h = self._size
        if out is None:
            out = np.empty((h, w, 4), dtype=np.uint8)
        self._check(self._L.gmlrt_download(self._h, out.ctypes.data, out.nbytes))
        return out

    def download_rows(self, out):
        """Only the rows this context rendered (its partition's bands) into the full-size raster `out` (uint8 [H,W,4])."""
        self._check(self._L.gmlrt_download_rows(self._h, out.ctypes.data, out.nbytes))
        return out

    def set_counting(self, enabled):
        self._check(self._L.gmlrt_set_counting(self._h, 1 if enabled else 0))

    def set_profiling(self, enabled):
        self._check(self._L.gmlrt_set_profiling(self._h, 1 if enabled else 0))

    def primary_hits(self):
        w, h = self._size
        ids = np.empty((h, w, 4), dtype=np.int32)
        self._check(self._L.gmlrt_primary_hits(self._h, ids.ctypes.data, ids.size))
        return ids

    def set_partition(self, part, n_parts):
        self._check(self._L.gmlrt_set_partition(self._h, part, n_parts))

    def ssim(self, a, b):
        """prim.SSIM (internal/prim/ssim.go:27-144) of two uint8 [H,W,4] rasters, computed on the GPU."""
        a = np.ascontiguousarray(a, dtype=np.uint8)
        b = np.ascontiguousarray(b, dtype=np.uint8)
        if a.shape != b.shape or a.ndim != 3 or a.shape[2] != 4:
            raise ValueError("images are not the same size")
        out = ctypes.c_double()
        self._check(self._L.gmlrt_ssim(self._h, a.ctypes.data, b.ctypes.data, a.shape[1], a.shape[0], ctypes.byref(out)))
        return out.value

    def ssim_device_noalloc(self, d_a, d_b, width, height):
        """gmlrt_ssim_device_noalloc: SSIM of two DEVICE rasters (addresses as ints), no allocation anywhere."""
        out = ctypes.c_double()
        self._check(self._L.gmlrt_ssim_device_noalloc(self._h, ctypes.c_void_p(d_a), ctypes.c_void_p(d_b), width, height, ctypes.byref(out)))
        return out.value

    def frame_ipc_export(self):
        """-> 64-byte CUDA IPC handle of this context's frame (after upload)."""
        buf = ctypes.create_string_buffer(IPC_HANDLE_BYTES)
        self._check(self._L.gmlrt_frame_ipc_export(self._h, buf, IPC_HANDLE_BYTES))
        return buf.raw

    def frame_ipc_attach(self, handle):
        """Store finished rows straight into another process's frame (NVLink peer mapping); None detaches."""
        if handle is None:
            self._check(self._L.gmlrt_frame_ipc_attach(self._h, None, 0))
        else:
            buf = ctypes.create_string_buffer(bytes(handle), IPC_HANDLE_BYTES)
            self._check(self._L.gmlrt_frame_ipc_attach(self._h, buf, IPC_HANDLE_BYTES))

    @property
    def frame_device_ptr(self):
        return self._L.gmlrt_frame_device(self._h)

    @property
    def stream(self):
        return self._L.gmlrt_stream(self._h)
