"""SURVEY §8 f3 on the GPU: material ids, texture coordinates and the RGB extension integrator, CUDA (through the C ABI) against
the CPU oracle, bit for bit.  Scenes come from the oracle's builder; every traversal-kernel variant renders them (gpu_ctx)."""
import numpy as np
import pytest

import materials_scenes as ms
import minipath_b200 as mp
from minipath_b200 import scenes
from oracle_lib import materials_table

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.fixture(scope="module")
def painted(oracle):
    s = ms.painted_room()
    mats = materials_table(s["albedos"], s["checker"], s["checker_scale"])
    return oracle.build_mesh(s["positions"], s["triangles"], s["normals"], texcoords=s["texcoords"], tri_material=s["tri_material"], materials=mats)


@pytest.fixture(scope="module")
def painted_camera():
    c = ms.PAINTED_CAMERA
    return mp.Camera.default().look_at(c["center"], c["target"], c["up"]).f_number(c["f_number"]).focus_distance(c["focus_distance"])


@pytest.fixture(scope="module")
def atrium_painted(oracle):
    """The Sponza stand-in with its nine per-part materials (minipath_b200/scenes.py: ATRIUM_PARTS)."""
    from oracle_lib import scenes

    p, n, t, uv, tm, table = scenes.atrium_mesh_materials()
    mats = materials_table(table[:, :3], table[:, 3] > 0, np.where(table[:, 3] > 0, table[:, 3], 1.0))
    return oracle.build_mesh(p, t, n, texcoords=uv, tri_material=tm, materials=mats)


def _setup(ctx, bvh, camera, w, h):
    ctx.upload_scene(bvh)
    s = camera.build_sampler((w, h))
    ctx.set_camera(s)
    ctx.create_film(w, h)
    return s


@pytest.mark.parametrize("scene_name", ["teapot", "cube", "painted"])
def test_hit_records_bit_exact(gpu_ctx, oracle, request, scene_name):
    """mp_intersect_rays: the full HitRecord (t, point, normal, material, texture_coords) of camera and incoherent rays."""
    bvh = request.getfixturevalue(scene_name)
    cam = request.getfixturevalue(scene_name + "_camera")
    w, h = 160, 96
    s = _setup(gpu_ctx, bvh, cam, w, h)
    st = mp.RenderSettings(64, 2, (w, h), seed=99)
    rays = oracle.generate_rays(s, w, mp.full_screen_tiles(w, h, 64), st.params())
    lo, hi = bvh.get_bounding_box()
    o, d = scenes.incoherent_rays(lo - 0.1, hi + 0.1, 8000, seed=5)
    rays = np.concatenate([rays, oracle.rays_from(o, d)])
    got = gpu_ctx.intersect_rays(rays)
    want = oracle.intersect_rays(bvh, rays)
    assert (want["hit"] == 1).sum() > 3000
    assert got.tobytes() == want.tobytes(), f"{(got.view(np.uint8).reshape(len(got), -1) != want.view(np.uint8).reshape(len(want), -1)).any(1).sum()} records differ"
    if scene_name == "painted":
        assert set(np.unique(want["material"][want["hit"] == 1])) == {0, 1, 2, 3}
        assert np.any(want["texture_coords"][:, :2] != 0)


@pytest.mark.parametrize("mode", [mp.MODE_MEGAKERNEL, mp.MODE_WAVEFRONT])
@pytest.mark.parametrize("bounces", [0, 1, 5])
def test_painted_room_film_bit_exact(gpu_ctx, oracle, painted, painted_camera, mode, bounces):
    """Four materials (two with the checker), flat and smooth normals: f32 film and the ray counts equal the oracle's."""
    if mode == mp.MODE_MEGAKERNEL and gpu_ctx.trace_variant not in (mp.TRACE_THREAD,):
        pytest.skip("the megakernel has one traversal; run once")
    w, h, spp = 200, 120, 3
    s = _setup(gpu_ctx, painted, painted_camera, w, h)
    st = mp.RenderSettings(64, spp, (w, h), max_bounces=bounces, seed=0xF3, mode=mode)
    tiles = mp.full_screen_tiles(w, h, 64, seed=3)
    gpu_ctx.enable_counters(True)
    gpu_ctx.reset_counters()
    gpu_ctx.clear_film()
    gpu_ctx.render_tiles(tiles, st.params())
    film = gpu_ctx.read_film()
    c = gpu_ctx.counters()
    gpu_ctx.enable_counters(False)
    ref, _, rc = oracle.render_tiles(painted, s, w, h, tiles, st.params())
    assert np.array_equal(_bits(film), _bits(ref)), f"{(_bits(film) != _bits(ref)).any(-1).sum()} pixels differ"
    assert (c["rays"], c["shadow_rays"], c["camera_rays"]) == (rc.rays, rc.shadow_rays, rc.camera_rays)
    if bounces:
        hit = ref[..., 3] > 0
        assert ((ref[..., 0] != ref[..., 1]) | (ref[..., 1] != ref[..., 2]))[hit].mean() > 0.9
    assert np.array_equal(gpu_ctx.resolve_rgba8(spp), oracle.resolve_rgba8(ref, spp))


def test_grey_table_gives_the_film_of_no_table(gpu_ctx, oracle, painted_camera):
    """The RGB kernels with an all-grey table against the scalar kernels without one: same bits (and both equal the oracle)."""
    s = ms.painted_room()
    plain = oracle.build_mesh(s["positions"], s["triangles"], s["normals"])
    grey = oracle.build_mesh(s["positions"], s["triangles"], s["normals"], texcoords=s["texcoords"], tri_material=s["tri_material"],
                             materials=materials_table(np.full((4, 3), 0.7, np.float32)))
    w, h = 160, 96
    st = mp.RenderSettings(64, 4, (w, h), max_bounces=6, seed=21, mode=mp.MODE_WAVEFRONT)
    tiles = mp.full_screen_tiles(w, h, 64)
    films = []
    for bvh in (plain, grey, plain):  # back to the scalar kernels on buffers that grew the RGB arrays
        sampler = _setup(gpu_ctx, bvh, painted_camera, w, h)
        gpu_ctx.clear_film()
        gpu_ctx.render_tiles(tiles, st.params())
        films.append(gpu_ctx.read_film())
    ref, _, _ = oracle.render_tiles(plain, sampler, w, h, tiles, st.params())
    for f in films:
        assert np.array_equal(_bits(f), _bits(ref))


def test_atrium_with_materials_bit_exact(gpu_ctx, oracle, atrium_painted, atrium_camera):
    """The stand-in with per-part materials, 5 bounces, two passes of samples."""
    w, h, spp = 160, 90, 2
    s = _setup(gpu_ctx, atrium_painted, atrium_camera, w, h)
    st = mp.RenderSettings(64, spp, (w, h), max_bounces=5, seed=31, mode=mp.MODE_WAVEFRONT)
    tiles = mp.full_screen_tiles(w, h, 64)
    gpu_ctx.clear_film()
    for base in (0, spp):
        p = st.params()
        p.sample_base = base
        gpu_ctx.render_tiles(tiles, p)
    film = gpu_ctx.read_film()
    ref = np.zeros((h, w, 4), np.float32)
    for base in (0, spp):
        p = st.params()
        p.sample_base = base
        oracle.render_tiles(atrium_painted, s, w, h, tiles, p, film=ref, threads=16)
    assert np.array_equal(_bits(film), _bits(ref)), f"{(_bits(film) != _bits(ref)).any(-1).sum()} pixels differ"
    assert ((ref[..., 0] != ref[..., 1]) | (ref[..., 1] != ref[..., 2])).mean() > 0.9


def test_render_api_with_materials(oracle, painted_camera):
    """mp_render (MODE_AUTO, host arrays in, RGBA8 out) on a product-built scene with the table equals the oracle's image."""
    s = ms.painted_room()
    mats = mp.make_materials(s["albedos"], s["checker"], s["checker_scale"])
    bvh = mp.TriangleBvh.from_mesh(s["positions"], s["triangles"], normals=s["normals"], texcoords=s["texcoords"], tri_material=s["tri_material"], materials=mats)
    w, h, spp = 192, 128, 4
    st = mp.RenderSettings(64, spp, (w, h), max_bounces=3, seed=5, mode=mp.MODE_AUTO)
    prog = mp.render(mp.Scene(bvh), painted_camera, st)
    prog.wait()
    img = np.array(prog.image())
    sampler = painted_camera.build_sampler((w, h))
    ref, _, _ = oracle.render_tiles(bvh, sampler, w, h, mp.full_screen_tiles(w, h, 64), st.params())
    assert np.array_equal(img, oracle.resolve_rgba8(ref, spp))
    assert np.any(img[..., 0] != img[..., 2])
