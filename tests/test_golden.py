"""Committed golden vectors (tests/golden/teapot_gate.npz, written by tests/golden/make_golden.py): the oracle must still
reproduce them bit for bit (CPU), and so must the CUDA path through the C ABI (GPU).  The fixture is an oracle regression pin —
the reference itself cannot be run here; see the generator's docstring."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import make_golden  # noqa: E402

import minipath_b200 as mp  # noqa: E402

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "teapot_gate.npz"))


def _same(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint8), np.ascontiguousarray(b).view(np.uint8))


def test_oracle_reproduces_golden(oracle):
    bvh, sampler, tiles = make_golden.case(mp)
    st = mp.RenderSettings(make_golden.TILE, 1, (make_golden.W, make_golden.H), max_bounces=0, seed=99)
    rays = oracle.generate_rays(sampler, make_golden.W, tiles, st.params())
    assert _same(rays, GOLD["rays"])
    assert _same(oracle.trace_rays(bvh, rays), GOLD["hits"])
    for bounces in (0, 5):
        stb = mp.RenderSettings(make_golden.TILE, 4, (make_golden.W, make_golden.H), max_bounces=bounces, seed=99)
        film, _, _ = oracle.render_tiles(bvh, sampler, make_golden.W, make_golden.H, tiles, stb.params())
        assert _same(film, GOLD[f"film_b{bounces}"])


@pytest.mark.gpu
def test_cuda_reproduces_golden(gpu_ctx):
    bvh, sampler, tiles = make_golden.case(mp)
    w, h = make_golden.W, make_golden.H
    gpu_ctx.upload_scene(bvh)
    gpu_ctx.set_camera(sampler)
    gpu_ctx.create_film(w, h)
    st = mp.RenderSettings(make_golden.TILE, 1, (w, h), max_bounces=0, seed=99)
    assert _same(gpu_ctx.generate_rays(tiles, st.params()), GOLD["rays"])
    assert _same(gpu_ctx.trace_rays(GOLD["rays"]), GOLD["hits"])
    for bounces, mode in ((0, mp.MODE_MEGAKERNEL), (0, mp.MODE_WAVEFRONT), (5, mp.MODE_MEGAKERNEL), (5, mp.MODE_WAVEFRONT)):
        stb = mp.RenderSettings(make_golden.TILE, 4, (w, h), max_bounces=bounces, seed=99, mode=mode)
        gpu_ctx.clear_film()
        gpu_ctx.render_tiles(tiles, stb.params())
        assert _same(gpu_ctx.read_film(), GOLD[f"film_b{bounces}"]), f"bounces={bounces} mode={mode}"
