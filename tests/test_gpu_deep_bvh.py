"""A BVH deeper than the 128-entry local stack of the thread-per-ray kernels (the reference split on a 30 k soup: depth 28, a 190-entry
stack): render() must pick a kernel that can run it — the pool kernel, whose stacks live in global scratch — and stay bit-exact; what
cannot run it fails with a message instead of overflowing."""
import ctypes as C

import numpy as np
import pytest

import minipath_b200 as mp
from minipath_b200 import _ffi, scenes

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def deep(oracle):
    p, t = scenes.soup_mesh(30000, seed=5)
    bvh = oracle.build_mesh(p, t)
    st = _ffi.MpBvhStats()  # the product's own (conservative: 7 * (depth - 1) + 1) bound decides which kernels may run the tree
    _ffi.check(_ffi.lib.mp_bvh_stats(bvh._ptr, C.byref(st)))
    assert 128 < st.max_stack <= 384, (st.max_depth, st.max_stack)
    return bvh


CAMERA = dict(center=(0.4, 0.3, 3.2), target=(0.0, 0.0, 0.0), up=(0.0, 1.0, 0.0))


def _camera():
    return mp.Camera.default().look_at(CAMERA["center"], CAMERA["target"], CAMERA["up"]).f_number(32.0).focus_distance(3.0)


@pytest.mark.parametrize("mode,bounces", [(mp.MODE_AUTO, 0), (mp.MODE_AUTO, 2), (mp.MODE_WAVEFRONT, 2)])
def test_render_on_a_deep_tree_is_bit_exact(oracle, deep, mode, bounces):
    w, h, spp = 128, 96, 2
    st = mp.RenderSettings(64, spp, (w, h), max_bounces=bounces, seed=9, mode=mode)
    prog = mp.render(mp.Scene(deep), _camera(), st)
    prog.wait()
    img = np.array(prog.image())
    sampler = _camera().build_sampler((w, h))
    ref, _, _ = oracle.render_tiles(deep, sampler, w, h, mp.full_screen_tiles(w, h, 64), st.params(), threads=16)
    assert (ref[..., 3] > 0).mean() > 0.3
    assert np.array_equal(img, oracle.resolve_rgba8(ref, spp))


def test_hits_on_a_deep_tree(oracle, deep):
    lo, hi = deep.get_bounding_box()
    o, d = scenes.incoherent_rays(lo - 0.1, hi + 0.1, 20000, seed=2)
    rays = oracle.rays_from(o, d)
    want = oracle.trace_rays(deep, rays, threads=16)
    for variant in (mp.TRACE_POOL, mp.TRACE_OCTET):
        ctx = mp.Context(0)
        ctx.set_trace_variant(variant)
        ctx.upload_scene(deep)
        assert ctx.trace_rays(rays).tobytes() == want.tobytes()
        ctx.close()
    ctx = mp.Context(0)  # thread-per-ray: refuses instead of overflowing its local stack
    ctx.upload_scene(deep)
    with pytest.raises(mp.MpError) as e:
        ctx.trace_rays(rays)
    assert "stack" in str(e.value)
    ctx.close()


def test_megakernel_refuses_a_deep_tree(deep):
    st = mp.RenderSettings(64, 1, (64, 64), max_bounces=0, seed=1, mode=mp.MODE_MEGAKERNEL)
    with pytest.raises(mp.MpError) as e:
        mp.render(mp.Scene(deep), _camera(), st).wait()
    assert "stack" in str(e.value)
