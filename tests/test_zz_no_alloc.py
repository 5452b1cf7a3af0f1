"""mp_trace_rays_device_noalloc (include/minipath_b200.h): the pool traversal on the library's static device scratch, no allocation
inside the call.  Checked through the C ABI against the existing entry point (mp_trace_rays_device, MP_TRACE_POOL) and the oracle,
bit for bit (the tolerance every hit-record test here uses), on two calls in a row — the second one finds the scratch as the first
left it.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import minipath_b200 as mp
from minipath_b200 import _ffi, scenes
from minipath_b200._abi_types import HIT_DTYPE, RAY_DTYPE

pytestmark = pytest.mark.gpu
lib = _ffi.lib


class _DevBuf:
    """Caller-owned device memory through the C ABI's raw helpers (the test may allocate; the entry point under test may not)."""

    def __init__(self, ctx, nbytes):
        self.ctx, self.ptr = ctx, C.c_void_p()
        _ffi.check(lib.mp_dev_alloc(ctx._h, nbytes, C.byref(self.ptr)))

    def upload(self, arr):
        _ffi.check(lib.mp_dev_upload(self.ctx._h, self.ptr, arr.ctypes.data, arr.nbytes))

    def download(self, arr):
        _ffi.check(lib.mp_dev_download(self.ctx._h, arr.ctypes.data, self.ptr, arr.nbytes))

    def fill(self, byte, nbytes):
        self.upload(np.full(nbytes, byte, np.uint8))

    def free(self):
        if self.ptr:
            lib.mp_dev_free(self.ctx._h, self.ptr)
            self.ptr = None


def _rays(oracle, bvh, camera, n_inc, seed):
    w, h = 192, 108
    s = camera.build_sampler((w, h))
    st = mp.RenderSettings(64, 2, (w, h), seed=seed)
    cam = oracle.generate_rays(s, w, mp.full_screen_tiles(w, h, 64), st.params())
    lo, hi = bvh.get_bounding_box()
    o, d = scenes.incoherent_rays(lo - 0.1, hi + 0.1, n_inc, seed=seed)
    return np.concatenate([cam, oracle.rays_from(o, d)])


@pytest.mark.parametrize("scene_name", ["teapot", "atrium"])
def test_noalloc_equals_the_allocating_entry_point_twice(oracle, request, scene_name):
    bvh = request.getfixturevalue(scene_name)
    camera = request.getfixturevalue(scene_name + "_camera")
    ctx = mp.Context(0)
    try:
        ctx.set_trace_variant(mp.TRACE_POOL)
        ctx.upload_scene(bvh)
        assert bvh.stats()["max_stack"] <= 128
        batches = [_rays(oracle, bvh, camera, 20000, seed=41), _rays(oracle, bvh, camera, 9000, seed=43)[::-1].copy()]
        n_max = max(len(b) for b in batches)
        d_rays, d_ref, d_new = _DevBuf(ctx, n_max * RAY_DTYPE.itemsize), _DevBuf(ctx, n_max * HIT_DTYPE.itemsize), _DevBuf(ctx, n_max * HIT_DTYPE.itemsize)
        for k, rays in enumerate(batches):  # two calls in a row on the shared static scratch, different sizes
            n = len(rays)
            d_rays.upload(rays)
            d_new.fill(0xAB, n * HIT_DTYPE.itemsize)  # a stale output buffer cannot pass for a result
            ctx.trace_rays_device(d_rays.ptr, n, d_ref.ptr)
            ctx.trace_rays_device_noalloc(d_rays.ptr, n, d_new.ptr)
            ctx.sync()
            ref, new = np.empty(n, HIT_DTYPE), np.empty(n, HIT_DTYPE)
            d_ref.download(ref), d_new.download(new)
            assert (ref["tri"] != mp.NO_HIT).sum() > 1000
            assert new.tobytes() == ref.tobytes(), f"call {k}: {(new.view(np.uint8).reshape(n, -1) != ref.view(np.uint8).reshape(n, -1)).any(1).sum()} of {n} hit records differ"
            assert new.tobytes() == oracle.trace_rays(bvh, rays, threads=16).tobytes()
        for b in (d_rays, d_ref, d_new):
            b.free()
    finally:
        ctx.close()


def test_noalloc_rejects_what_it_cannot_serve(teapot):
    ctx = mp.Context(0)
    try:
        with pytest.raises(mp.MpError) as e:  # no scene yet
            ctx.trace_rays_device_noalloc(C.c_void_p(16), 1, C.c_void_p(16))
        assert e.value.code == _ffi.MP_ESTATE
        ctx.upload_scene(teapot)
        with pytest.raises(mp.MpError) as e:  # more rays than the stated maximum: refused before anything is touched
            ctx.trace_rays_device_noalloc(C.c_void_p(16), (1 << 30) + 1, C.c_void_p(16))
        assert e.value.code == _ffi.MP_EINVAL and "at most" in str(e.value)
        with pytest.raises(mp.MpError):
            ctx.trace_rays_device_noalloc(None, 5, None)
        ctx.trace_rays_device_noalloc(None, 0, None)  # nothing to do is not an error
    finally:
        ctx.close()
