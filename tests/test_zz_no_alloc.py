"""gmlrt_ssim_device_noalloc (include/gmlrt.h): the allocation-free variant of gmlrt_ssim, whose per-block sums live in
a file-scope __device__ array.  Collected last on purpose.  The check: through the
C ABI, on device rasters, the result equals gmlrt_ssim's on the same pixels - on two calls in a row that use
different amounts of the shared scratch, and again with the operands swapped."""
import numpy as np
import pytest

from go_raytracer_b200 import scenes

TOL = 1e-12          # the tolerance of test_gpu_ssim_matches_oracle; the two entry points sum in the same order


def _rendered(cuda, precision, blob):
    ctx = cuda.Context((0,), precision)
    ctx.upload(blob)
    ctx.trace()
    return ctx, ctx.download()


@pytest.mark.gpu
def test_noalloc_ssim_equals_gmlrt_ssim_on_two_calls_in_a_row():
    from go_raytracer_b200 import cuda
    big = scenes.to_blob(scenes.mixed(320, 200, grid=(8, 4, 4), cones=False, depth=3))
    small = scenes.to_blob(scenes.spheres_lattice(96, 64, grid=(4, 2, 2)))
    for blob in (big, small):                          # the second call uses a prefix of the scratch the first one filled
        w, h = cuda.Context._blob_size(blob)
        a, img_a = _rendered(cuda, cuda.PRECISION_F32, blob)
        b, img_b = _rendered(cuda, cuda.PRECISION_F64_PARITY, blob)
        try:
            want = a.ssim(img_a, img_b)                # existing entry point: host rasters, cudaMalloc'ed scratch
            assert 0.9 < want <= 1.0
            got = a.ssim_device_noalloc(a.frame_device_ptr, b.frame_device_ptr, w, h)
            assert abs(got - want) <= TOL, (got, want)
            again = a.ssim_device_noalloc(b.frame_device_ptr, a.frame_device_ptr, w, h)
            assert abs(again - a.ssim(img_b, img_a)) <= TOL
            assert abs(a.ssim_device_noalloc(a.frame_device_ptr, a.frame_device_ptr, w, h) - 1.0) <= TOL
        finally:
            a.close()
            b.close()


@pytest.mark.gpu
def test_noalloc_ssim_rejects_sizes_beyond_the_static_scratch():
    from go_raytracer_b200 import cuda
    blob = scenes.to_blob(scenes.spheres_lattice(96, 64, grid=(4, 2, 2)))
    ctx, _ = _rendered(cuda, cuda.PRECISION_F32, blob)
    try:
        p = ctx.frame_device_ptr
        with pytest.raises(cuda.GmlrtError) as ei:     # 8192 x 8192 > 2^25 windows; rejected before anything is launched
            ctx.ssim_device_noalloc(p, p, 8192, 8192)
        assert ei.value.code == -1 and "GMLRT_SSIM_NOALLOC_MAX_WINDOWS" in str(ei.value)
        with pytest.raises(cuda.GmlrtError):
            ctx.ssim_device_noalloc(p, p, 10, 64)      # smaller than the 11x11 window, like gmlrt_ssim
        assert abs(ctx.ssim_device_noalloc(p, p, 96, 64) - 1.0) <= TOL
    finally:
        ctx.close()
