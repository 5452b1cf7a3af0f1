"""mp_group_* and mp_render on several devices: the multi-GPU half of the C ABI (reference scheduler
src/renderer/machinery.rs:51-116,205-209).  Tile sharding must reproduce the single-device film bit for bit (every pixel is
written by exactly one device, the NCCL sum adds zeros); the spp split must equal the sum of the per-device partial films."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W, H, TS = 320, 192, 32


def _settings(mp, spp, bounces, seed=11):
    return mp.RenderSettings(TS, spp, (W, H), max_bounces=bounces, seed=seed, mode=mp.MODE_WAVEFRONT)


def _single(mp, bvh, sampler, tiles, st, sample_base=0, sample_count=None, variant=None):
    ctx = mp.Context(0)
    ctx.set_trace_variant(mp.TRACE_HYBRID_POOL_S0 if variant is None else variant)
    ctx.upload_scene(bvh)
    ctx.set_camera(sampler)
    ctx.create_film(W, H)
    ctx.render_tiles(tiles, st.params(sample_count=sample_count, sample_base=sample_base))
    film = ctx.read_film()
    img = ctx.resolve_rgba8(st.sample_count)
    ctx.close()
    return film, img


def _need(n):
    import minipath_b200 as mp

    if mp.device_count() < n:
        pytest.skip(f"needs {n} CUDA devices, have {mp.device_count()}")


@pytest.mark.parametrize("shard", ["tiles", "spp"])
@pytest.mark.parametrize("bounces", [0, 3])
def test_group_of_one_device_matches_the_context(teapot, teapot_camera, shard, bounces):
    import minipath_b200 as mp

    sampler = teapot_camera.build_sampler((W, H))
    tiles = mp.full_screen_tiles(W, H, TS)
    st = _settings(mp, 6, bounces)
    want_film, want_img = _single(mp, teapot, sampler, tiles, st)
    g = mp.Group(devices=[0])
    assert g.size() == (1, 1)
    g.set_trace_variant(mp.TRACE_HYBRID_POOL_S0)
    g.upload_scene(teapot)
    g.set_camera(sampler)
    g.create_film(W, H)
    for _ in range(2):  # the second frame exercises the rewind of the tile counter and the film clear
        g.clear_film()
        g.render(tiles, st.params(), shard=mp.SHARD_TILES if shard == "tiles" else mp.SHARD_SPP, min_batch=3)
        g.reduce_film()
        img = g.resolve_rgba8(st.sample_count)
        film = g.read_film(0)
        assert np.array_equal(film.view(np.uint32), want_film.view(np.uint32))
        assert np.array_equal(img, want_img)
    stats = g.member_stats(0)
    assert stats["tiles"] == len(tiles) and stats["batches"] >= (1 if shard == "spp" else 2)
    assert g.last_ms() > 0.0
    g.close()


def test_group_tile_sharding_two_devices_bit_exact(teapot, teapot_camera):
    _need(2)
    import minipath_b200 as mp

    sampler = teapot_camera.build_sampler((W, H))
    tiles = mp.full_screen_tiles(W, H, TS)
    st = _settings(mp, 8, 4)
    want_film, want_img = _single(mp, teapot, sampler, tiles, st)
    g = mp.Group(devices=[0, 1])
    g.set_trace_variant(mp.TRACE_HYBRID_POOL_S0)
    g.upload_scene(teapot)
    g.set_camera(sampler)
    g.create_film(W, H)
    g.render(tiles, st.params(), shard=mp.SHARD_TILES, min_batch=2)
    parts = [g.read_film(i) for i in range(2)]
    covered = [(p[..., 3] != 0) | (p[..., 0] != 0) for p in parts]
    assert not np.any(covered[0] & covered[1]), "a pixel was rendered by both devices"
    s0, s1 = g.member_stats(0), g.member_stats(1)
    assert s0["tiles"] + s1["tiles"] == len(tiles) and s0["tiles"] > 0 and s1["tiles"] > 0
    g.reduce_film()
    img = g.resolve_rgba8(st.sample_count)
    film = g.read_film(0)
    assert np.array_equal(film.view(np.uint32), want_film.view(np.uint32))
    assert np.array_equal(img, want_img)
    g.close()


def test_group_spp_split_two_devices_is_the_sum_of_the_partial_films(teapot, teapot_camera):
    _need(2)
    import minipath_b200 as mp

    sampler = teapot_camera.build_sampler((W, H))
    tiles = mp.full_screen_tiles(W, H, TS)
    st = _settings(mp, 7, 3)  # 7 = 4 + 3: the remainder goes to the low rank
    a, _ = _single(mp, teapot, sampler, tiles, st, sample_base=0, sample_count=4)
    b, _ = _single(mp, teapot, sampler, tiles, st, sample_base=4, sample_count=3)
    g = mp.Group(devices=[0, 1])
    g.set_trace_variant(mp.TRACE_HYBRID_POOL_S0)
    g.upload_scene(teapot)
    g.set_camera(sampler)
    g.create_film(W, H)
    g.render(tiles, st.params(), shard=mp.SHARD_SPP)
    assert np.array_equal(g.read_film(0).view(np.uint32), a.view(np.uint32))
    assert np.array_equal(g.read_film(1).view(np.uint32), b.view(np.uint32))
    g.reduce_film()
    g.sync()
    assert np.array_equal(g.read_film(0).view(np.uint32), (a + b).view(np.uint32))  # two addends: the sum has one order
    g.close()


def test_mp_render_on_two_devices_equals_one_device(teapot, teapot_camera):
    _need(2)
    import minipath_b200 as mp

    st = mp.RenderSettings(TS, 6, (W, H), max_bounces=2, seed=5)
    one = mp.render(mp.Scene(teapot), teapot_camera, st, devices=[0])
    one.wait()
    finished = []
    two = mp.render(mp.Scene(teapot), teapot_camera, st, finished_tile_callback=lambda t, s: finished.append((t, s.finished)), devices=[0, 1])
    two.wait()
    assert np.array_equal(two.image(), one.image())
    assert len(finished) == len(mp.full_screen_tiles(W, H, TS)) and two.progress().finished == len(finished)


def test_mp_render_progress_and_abort_are_batch_granular(teapot, teapot_camera):
    """A long single-device render is cut into batches of about one wavefront pass (host_render.cpp): progress() advances and
    abort() takes effect before the frame is complete (machinery.rs:133-142,161-165)."""
    import minipath_b200 as mp

    st = mp.RenderSettings(64, 512, (1280, 768), max_bounces=4, seed=3)  # 240 tiles x 2 M paths each: 15 batches of 16 tiles
    seen = []
    prog = mp.render(mp.Scene(teapot), teapot_camera, st, finished_tile_callback=lambda t, s: seen.append(s.finished), devices=[0])
    import time

    t0 = time.time()
    while not seen and time.time() - t0 < 60:
        time.sleep(0.002)
    prog.abort()
    prog.wait()
    total = prog.progress().total
    assert 0 < prog.progress().finished < total, (prog.progress().finished, total)
    assert sorted(seen) == seen
