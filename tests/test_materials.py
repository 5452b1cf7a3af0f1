"""SURVEY §8 f3 on the CPU: the .mtl loaders, the material id and texture coordinate carried through hit finalisation
(geometry/mod.rs:70-80; ray_bvh_intersection.rs:69-94; triangle.rs:229-244) and the oracle's RGB integrator."""
import os

import numpy as np
import pytest

import materials_scenes as ms
from oracle_lib import _ffi, materials_table

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SPONZA_MTL = os.path.join(ROOT, "data", "Sponza", "sponza.mtl")

OBJ = """mtllib paint.mtl
v -1 0 -1
v 1 0 -1
v 1 0 1
v -1 0 1
v 0 1 0
vt 0 0
vt 1 0
vt 1 1
vt 0 1
vt 0.5 0.5
f 1/1 3/3 2/2
usemtl brick
f 1/1 4/4 3/3
usemtl glaze
f 1/1 2/2 5/5
f 2/2 3/3 5/5
usemtl brick
f 3/3 4/4 5/5
"""
MTL = """# three entries, one without Kd, one with a diffuse map
newmtl brick
Kd 0.6 0.25 0.125
newmtl plain
newmtl glaze
Kd 0.1 0.5 0.75
map_Kd glaze.png
"""


@pytest.fixture(scope="module")
def painted(oracle):
    s = ms.painted_room()
    mats = materials_table(s["albedos"], s["checker"], s["checker_scale"])
    return oracle.build_mesh(s["positions"], s["triangles"], s["normals"], texcoords=s["texcoords"], tri_material=s["tri_material"], materials=mats)


def _write_obj(tmp_path):
    (tmp_path / "paint.mtl").write_text(MTL)
    p = tmp_path / "paint.obj"
    p.write_text(OBJ)
    return str(p)


def _material_of_source_triangle(bvh):
    """material id per real triangle, keyed by its sorted vertex ids (the builders reorder triangles)."""
    sh = bvh.tri_shading()
    real = ~((sh[:, 0] == 0) & (sh[:, 1] == 0) & (sh[:, 2] == 0))
    return {tuple(sorted(int(x) for x in r[:3])): int(r[3] & _ffi.TRI_MATERIAL_MASK) for r in sh[real]}


def test_mtl_loaders_agree_and_assign_the_ids(oracle, tmp_path):
    import minipath_b200 as mp

    path = _write_obj(tmp_path)
    a, b = oracle.build_obj(path), mp.TriangleBvh.with_obj(path)
    for bvh in (a, b):
        m = bvh.materials()
        assert len(m) == 4  # default + brick, plain, glaze in file order
        assert np.allclose(m["albedo"], [[0.7, 0.7, 0.7], [0.6, 0.25, 0.125], [0.7, 0.7, 0.7], [0.1, 0.5, 0.75]])
        assert list(m["flags"]) == [0, 0, 0, _ffi.MAT_CHECKER]
        assert np.all(m["checker_scale"] == np.float32(4.0))
        ids = _material_of_source_triangle(bvh)
        # vertex ids = first appearance of (v, vt): 1→0, 3→1, 2→2, 4→3, 5→4
        assert ids == {(0, 1, 2): 0, (0, 1, 3): 1, (0, 2, 4): 3, (1, 2, 4): 3, (1, 3, 4): 1}
    assert a.materials().tobytes() == b.materials().tobytes()
    assert a.tri_shading().tobytes() == b.tri_shading().tobytes()
    assert a.vertices().tobytes() == b.vertices().tobytes()


def test_obj_without_mtllib_has_no_table(teapot, cube):
    assert teapot.raw.n_materials == 0 and cube.raw.n_materials == 0
    assert not np.any(teapot.tri_shading()[:, 3] & _ffi.TRI_MATERIAL_MASK)


def test_unknown_material_and_missing_library_are_errors(oracle, tmp_path):
    import minipath_b200 as mp

    (tmp_path / "a.obj").write_text("mtllib nowhere.mtl\nv 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n")
    (tmp_path / "b.mtl").write_text("newmtl x\nKd 1 1 1\n")
    (tmp_path / "b.obj").write_text("mtllib b.mtl\nv 0 0 0\nv 1 0 0\nv 0 1 0\nusemtl y\nf 1 2 3\n")
    for name in ("a.obj", "b.obj"):
        with pytest.raises(RuntimeError):
            oracle.build_obj(str(tmp_path / name))
        with pytest.raises(mp.MpError):
            mp.TriangleBvh.with_obj(str(tmp_path / name))
    with pytest.raises(mp.MpError):  # id outside the table
        mp.TriangleBvh.from_mesh([[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[0, 1, 2]], tri_material=[2], materials=mp.make_materials([[1, 1, 1], [0, 0, 0]]))
    with pytest.raises(RuntimeError):
        oracle.build_mesh([[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[0, 1, 2]], tri_material=[1])


@pytest.mark.skipif(not os.path.exists(SPONZA_MTL), reason="data/Sponza/sponza.mtl not present")
def test_the_reference_sponza_mtl_parses(oracle, tmp_path):
    """The one material library the reference ships (data/Sponza/sponza.mtl): every newmtl becomes an entry with its Kd, and the
    entries with a map_Kd get the procedural checker (the images are not in the data set)."""
    import minipath_b200 as mp

    names, kd, mapped = [], {}, set()
    for line in open(SPONZA_MTL):
        w = line.split()
        if not w:
            continue
        if w[0] == "newmtl":
            names.append(w[1])
        elif w[0] == "Kd":
            kd[names[-1]] = [float(x) for x in w[1:4]]
        elif w[0] == "map_Kd":
            mapped.add(names[-1])
    assert len(names) >= 20
    obj = tmp_path / "s.obj"
    obj.write_text(f"mtllib {SPONZA_MTL}\nv 0 0 0\nv 1 0 0\nv 0 1 0\nusemtl {names[3]}\nf 1 2 3\n")
    for bvh in (oracle.build_obj(str(obj)), mp.TriangleBvh.with_obj(str(obj))):
        m = bvh.materials()
        assert len(m) == len(names) + 1
        for k, n in enumerate(names):
            assert np.allclose(m["albedo"][k + 1], kd.get(n, [0.7] * 3))
            assert bool(m["flags"][k + 1] & _ffi.MAT_CHECKER) == (n in mapped)
        assert int(bvh.tri_shading()[0, 3] & _ffi.TRI_MATERIAL_MASK) == 4


def test_hit_record_carries_material_and_texture_coordinates(oracle, painted):
    """Rays straight down onto the floor (material 0, uv = 2 * (x+4, z+4) / 8 ... per the grid) and onto the slab (material 2)."""
    s = ms.painted_room()
    xs = np.array([[-3.3, -3.1], [0.4, 3.7], [3.9, -0.2], [-0.5, 0.0]], np.float32)
    rays = oracle.rays_from([(x, 6.0, z) for x, z in xs], [(0.0, -1.0, 0.0)] * len(xs))
    rec = oracle.intersect_rays(painted, rays)
    assert list(rec["hit"]) == [1, 1, 1, 1]
    assert list(rec["material"][:3]) == [0, 0, 0] and rec["material"][3] == 2  # the last one lands on the slab
    # floor grid: u runs along +z, v along +x (materials_scenes.painted_room), uv_scale 2 → uv = 2 * ((z+4)/8, (x+4)/8)
    want = 2.0 * np.stack([(xs[:3, 1] + 4) / 8, (xs[:3, 0] + 4) / 8], -1)
    assert np.allclose(rec["texture_coords"][:3, :2], want, atol=2e-3)  # positions are quantised to 16 bits, uv are not
    assert np.all(rec["texture_coords"][:, 2] == 0)
    assert np.allclose(rec["normal"][:3], [[0, 1, 0]] * 3, atol=1e-3) or np.allclose(rec["normal"][:3], [[0, -1, 0]] * 3, atol=1e-3)
    assert s["tri_material"].max() == 3


def _render(oracle, bvh, sampler, w, h, spp, bounces, seed=7):
    p = _ffi.MpRenderParams(spp, 0, bounces, 1, 0, 64, seed)
    film, _, c = oracle.render_tiles(bvh, sampler, w, h, oracle.tiles(w, h, 64), p)
    return film, c


def test_grey_table_gives_the_film_of_no_table(oracle):
    s = ms.painted_room()
    cam = ms.PAINTED_CAMERA
    sampler = oracle.camera_sampler(cam["center"], cam["target"], cam["up"], (96, 64), f_number=cam["f_number"], focus_distance=cam["focus_distance"])
    plain = oracle.build_mesh(s["positions"], s["triangles"], s["normals"])
    grey = oracle.build_mesh(s["positions"], s["triangles"], s["normals"], texcoords=s["texcoords"], tri_material=s["tri_material"],
                             materials=materials_table(np.full((4, 3), 0.7, np.float32)))
    for bounces in (0, 1, 6):
        a, ca = _render(oracle, plain, sampler, 96, 64, 4, bounces)
        b, cb = _render(oracle, grey, sampler, 96, 64, 4, bounces)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
        assert ca.rays == cb.rays and ca.shadow_rays == cb.shadow_rays


def test_colour_and_the_reference_shader(oracle, painted):
    cam = ms.PAINTED_CAMERA
    sampler = oracle.camera_sampler(cam["center"], cam["target"], cam["up"], (96, 64), f_number=cam["f_number"], focus_distance=cam["focus_distance"])
    film0, _ = _render(oracle, painted, sampler, 96, 64, 2, 0)
    assert np.array_equal(film0[..., 0], film0[..., 1]) and np.array_equal(film0[..., 1], film0[..., 2])  # worker.rs:57-64 ignores materials
    film, _ = _render(oracle, painted, sampler, 96, 64, 8, 4)
    hit = film[..., 3] > 0
    assert hit.sum() > 3000
    differs = (film[..., 0] != film[..., 1]) | (film[..., 1] != film[..., 2])
    assert differs[hit].mean() > 0.9
    wall = film[8:20, 30:60]  # the red back wall fills the upper middle of the frame
    assert wall[..., 0].mean() > 2.0 * wall[..., 2].mean()


def test_checker_known_answer(oracle):
    """One square floor, sun above, camera straight down, one bounce: every path is floor hit → unoccluded sun ray → bounce into the
    sky, so a pixel whose samples stay inside one checker cell holds  albedo(cell) * (cos_sun + 1)  per channel, exactly the
    same float for every sample."""
    pos, uv, tri = ms.single_floor(8.0)
    albedo = np.array([0.8, 0.4, 0.2], np.float32)
    bvh = oracle.build_mesh(pos, tri, texcoords=uv, tri_material=[0, 0], materials=materials_table([albedo], [True], [1.0]))
    w = h = 64
    sampler = oracle.camera_sampler((0.0, 10.0, 0.0), (0.0, 0.0, 0.0), (0.0, 0.0, -1.0), (w, h), f_number=1e6)
    spp = 4
    film, c = _render(oracle, bvh, sampler, w, h, spp, 1)
    assert c.shadow_rays == c.camera_rays  # every camera ray hit the floor and faced the sun
    cos_sun = np.float32(0.86386842)
    full = albedo * cos_sun + albedo  # (T * E) * c + T * ENV with E = ENV = 1
    mean = film[..., :3] / spp
    is_full = np.all(np.abs(mean - full) < 1e-6, axis=-1)
    is_half = np.all(np.abs(mean - 0.5 * full) < 1e-6, axis=-1)
    assert (is_full | is_half).mean() > 0.7  # the rest straddle a cell border
    assert 0.25 < is_full.sum() / max(1, (is_full | is_half).sum()) < 0.75
    # the pattern is a checkerboard in world space: recover the cell of each clean pixel from a primary hit record
    p = _ffi.MpRenderParams(1, 0, 0, 1, 0, 64, 7)
    rays = oracle.generate_rays(sampler, w, oracle.tiles(w, h, 64), p)
    rec = oracle.intersect_rays(bvh, rays)
    cell = np.floor(rec["texture_coords"][:, 0]) + np.floor(rec["texture_coords"][:, 1])
    odd = (cell.astype(np.int64) & 1).reshape(h, w).astype(bool)
    clean = is_full | is_half
    inner = clean & (np.abs(rec["texture_coords"][:, 0] % 1 - 0.5).reshape(h, w) < 0.3) & (np.abs(rec["texture_coords"][:, 1] % 1 - 0.5).reshape(h, w) < 0.3)
    assert inner.sum() > 500
    assert np.array_equal(is_half[inner], odd[inner])
