"""the device BVH8 builder (mp_bvh_build_device, csrc/build_device.cu).

Its split is not the reference's, so its arrays cannot be compared with the host builders'; what is checked instead:
the reference FORMAT's invariants against the source mesh (tests/bvh_checks.py, the same checker the host builders pass), the
oracle's traversal of the tree against a brute force over all its packets and against the CUDA traversal (bit for bit), the film,
and the tree's quality as visits per ray next to the host reference-split tree."""
import os

import numpy as np
import pytest

import minipath_b200 as mp
from bvh_checks import check_builder_invariants
from minipath_b200 import scenes
from test_independent import DATA, parse_obj

pytestmark = pytest.mark.gpu
DEVICE = mp.TriangleBvh.BUILDER_DEVICE


def _mesh(name):
    if name in ("teapot", "cube"):
        pos, tri = parse_obj(os.path.join(DATA, name + ".obj"), int(name == "cube"))
        return pos, tri, None
    if name == "atrium":
        p, n, t = scenes.atrium_mesh()
        return p, t, n
    if name == "soup":
        p, t = scenes.soup_mesh(100_000)
        return p, t, None
    rng = np.random.default_rng(5)
    if name == "one-leaf":  # no more triangles than the builder's leaf size: the root is a leaf link (building.rs:109-113)
        p = rng.uniform(-1, 1, (3 * 7, 3)).astype(np.float32)
        return p, np.arange(21, dtype=np.uint32).reshape(-1, 3), None
    if name == "forty":
        p = rng.uniform(-1, 1, (3 * 40, 3)).astype(np.float32)
        return p, np.arange(120, dtype=np.uint32).reshape(-1, 3), None
    if name == "57":
        p = rng.uniform(-1, 1, (3 * 57, 3)).astype(np.float32)
        return p, np.arange(171, dtype=np.uint32).reshape(-1, 3), None
    if name == "stacked":  # 700 triangles with ONE centroid: no Morton bit separates them, the split falls back to the middle
        base = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
        k = 700
        ang = rng.uniform(0, 2 * np.pi, k)
        c, s = np.cos(ang), np.sin(ang)
        ctr = base.mean(0)
        q = base - ctr
        p = np.stack([np.stack([c[:, None] * q[None, :, 0] - s[:, None] * q[None, :, 1], s[:, None] * q[None, :, 0] + c[:, None] * q[None, :, 1],
                                np.broadcast_to(q[None, :, 2], (k, 3))], -1)], 0)[0] + ctr
        return p.reshape(-1, 3).astype(np.float32), np.arange(3 * k, dtype=np.uint32).reshape(-1, 3), None
    raise KeyError(name)


_CACHE = {}


def _built(name):
    if name not in _CACHE:
        pos, tri, nrm = _mesh(name)
        _CACHE[name] = (pos, tri, nrm, mp.TriangleBvh.from_mesh(pos, tri, normals=nrm, builder=DEVICE))
    return _CACHE[name]


def _rays(oracle, bvh, n_cam, n_inc, seed):
    lo, hi = bvh.get_bounding_box()
    ctr, s = 0.5 * (lo + hi), float(np.linalg.norm(hi - lo))
    w = h = int(np.sqrt(n_cam))
    sampler = oracle.camera_sampler(ctr + np.array([0.3, 0.4, 1.0]) * s, ctr, (0, 1, 0), (w, h), f_number=64.0, focus_distance=s)
    st = mp.RenderSettings(64, 1, (w, h), seed=seed)
    cam = oracle.generate_rays(sampler, w, mp.full_screen_tiles(w, h, 64), st.params())
    o, d = scenes.incoherent_rays(lo - 0.05 * s, hi + 0.05 * s, n_inc, seed=seed)
    return np.concatenate([cam, oracle.rays_from(o, d)])


@pytest.mark.parametrize("name", ["teapot", "cube", "atrium", "soup", "one-leaf", "forty", "57", "stacked"])
def test_device_tree_keeps_the_format_invariants(name):
    pos, tri, _, bvh = _built(name)
    check_builder_invariants(pos, tri, bvh)
    st = bvh.stats()
    assert st["n_triangles"] == len(tri) and st["max_stack"] <= 128
    if name == "one-leaf":
        assert bvh.raw.n_inner == 0 and (bvh.raw.root & 7) == 1
    print(name, st, "device build ms", bvh.last_build_ms)


@pytest.mark.parametrize("name", ["teapot", "cube", "atrium", "soup", "forty", "57", "stacked"])
def test_oracle_and_cuda_traverse_the_device_tree_bit_equal(gpu_ctx, oracle, name):
    _, _, _, bvh = _built(name)
    rays = _rays(oracle, bvh, 128 * 128, 12000, seed=11)
    gpu_ctx.upload_scene(bvh)
    got = gpu_ctx.trace_rays(rays)
    want = oracle.trace_rays(bvh, rays, threads=16)
    assert (want["tri"] != mp.NO_HIT).sum() > 500
    assert got.tobytes() == want.tobytes(), f"{(got.view(np.uint8).reshape(len(got), -1) != want.view(np.uint8).reshape(len(want), -1)).any(1).sum()} rays differ"


@pytest.mark.parametrize("name", ["teapot", "atrium", "soup"])
def test_device_tree_hides_nothing(oracle, name):
    """Traversal of the device-built tree == brute force over all its packets (closest t bit for bit).  Not run on "stacked": its
    700 triangles are coplanar, every box has zero thickness, and the reference's `t1 > best.t` cull (ray_bvh_intersection.rs:40) may
    then drop a leaf whose copy of the plane is one ulp nearer — a property of the traversal rule on degenerate input, not of a builder."""
    _, _, _, bvh = _built(name)
    rays = _rays(oracle, bvh, 64 * 64, 4000 if name != "teapot" else 12000, seed=3)
    fast, slow = oracle.trace_rays(bvh, rays, threads=16), oracle.brute_force(bvh, rays, threads=16)
    assert np.array_equal(fast["tri"] == mp.NO_HIT, slow["tri"] == mp.NO_HIT)
    hit = slow["tri"] != mp.NO_HIT
    assert np.array_equal(fast["t"][hit].view(np.uint32), slow["t"][hit].view(np.uint32))


def test_film_on_the_device_tree_bit_exact(gpu_ctx, oracle, atrium_camera):
    _, _, _, bvh = _built("atrium")
    w, h, spp = 160, 90, 2
    gpu_ctx.upload_scene(bvh)
    s = atrium_camera.build_sampler((w, h))
    gpu_ctx.set_camera(s)
    gpu_ctx.create_film(w, h)
    st = mp.RenderSettings(64, spp, (w, h), max_bounces=4, seed=17, mode=mp.MODE_WAVEFRONT)
    tiles = mp.full_screen_tiles(w, h, 64)
    gpu_ctx.clear_film()
    gpu_ctx.render_tiles(tiles, st.params())
    ref, _, _ = oracle.render_tiles(bvh, s, w, h, tiles, st.params(), threads=16)
    assert np.array_equal(gpu_ctx.read_film().view(np.uint32), ref.view(np.uint32))


def test_build_is_deterministic_and_carries_materials():
    p, n, t, uv, tm, table = scenes.atrium_mesh_materials()
    mats = mp.make_materials(table[:, :3], checker=table[:, 3] > 0, checker_scale=np.where(table[:, 3] > 0, table[:, 3], 1.0))
    a = mp.TriangleBvh.from_mesh(p, t, normals=n, builder=DEVICE, texcoords=uv, tri_material=tm, materials=mats)
    b = mp.TriangleBvh.from_mesh(p, t, normals=n, builder=DEVICE, texcoords=uv, tri_material=tm, materials=mats)
    for f in ("inner_bounds", "inner_links", "inner_boxes", "packets", "packet_boxes", "tri_shading", "vertices"):
        assert getattr(a, f)().tobytes() == getattr(b, f)().tobytes(), f
    assert a.materials().tobytes() == mats.tobytes()
    sh = a.tri_shading()
    real = sh[:, :3].any(axis=1)
    key = lambda v: np.sort(v.astype(np.uint64), axis=1) @ np.array([1, 1 << 21, 1 << 42], np.uint64)  # noqa: E731
    want = dict(zip(key(t).tolist(), tm.tolist()))
    got = dict(zip(key(sh[real, :3]).tolist(), (sh[real, 3] & 0x7FFFFFFF).tolist()))
    assert got == want
    host = mp.TriangleBvh.from_mesh(p, t, normals=n, builder=0, texcoords=uv, tri_material=tm, materials=mats)
    assert a.vertices().tobytes() == host.vertices().tobytes()  # the vertex records do not depend on the builder


def test_bad_input_is_rejected():
    tri = np.array([[0, 1, 2]], np.uint32)
    pos = np.zeros((3, 3), np.float32)
    with pytest.raises(mp.MpError):
        mp.TriangleBvh.from_mesh(pos, np.zeros((0, 3), np.uint32), builder=DEVICE)
    with pytest.raises(mp.MpError):
        mp.TriangleBvh.from_mesh(pos, np.array([[0, 1, 3]], np.uint32), builder=DEVICE)
    with pytest.raises(mp.MpError):
        mp.TriangleBvh.from_mesh(pos, tri, builder=DEVICE, tri_material=[1])
    with pytest.raises(mp.MpError):
        mp.TriangleBvh.from_mesh(pos, tri, builder=DEVICE, device=99)


QUALITY_FACTOR = 1.25  # sectors gathered per ray on the device-built tree, at most this multiple of the host reference-split tree's


def _visits(ctx, bvh, rays):
    ctx.upload_scene(bvh)
    ctx.enable_counters(True)
    ctx.reset_counters()
    hits = ctx.trace_rays(rays)
    c = ctx.counters()
    ctx.enable_counters(False)
    st = bvh.stats()
    return {"inner": round(c["inner_visits"] / len(rays), 2), "packets": round(c["packet_visits"] / len(rays), 2), "hits": int((hits["tri"] != mp.NO_HIT).sum()),
            "sectors": round((5 * c["inner_visits"] + 6 * c["packet_visits"]) / len(rays), 2), "n_inner": st["n_inner"], "n_leaves": st["n_leaves"],
            "n_packets": st["n_packets"], "leaf_fill": round(st["mean_leaf_fill"], 3)}


@pytest.mark.parametrize("name", ["teapot", "atrium"])
def test_quality_next_to_the_reference_split(oracle, name, monkeypatch):
    """Visits per ray (camera + incoherent rays) on the device-built tree and on the host reference-split tree, as L1 sectors
    (5 per node record, 6 per packet record: SURVEY §8(d)); the leaf-size knob of the device builder is swept beside the default."""
    pos, tri, nrm, dev = _built(name)
    host = mp.TriangleBvh.from_mesh(pos, tri, normals=nrm, builder=0)
    rays = _rays(oracle, dev, 256 * 256, 60000, seed=23)
    ctx = mp.Context(0)
    ctx.set_trace_variant(mp.TRACE_POOL)
    out = {"host reference split": _visits(ctx, host, rays), "device (default leaf size)": _visits(ctx, dev, rays)}
    for leaf in (8, 16, 24, 32, 56):
        monkeypatch.setenv("MINIPATH_B200_DEVICE_LEAF", str(leaf))
        out[f"device, leaf <= {leaf}"] = _visits(ctx, mp.TriangleBvh.from_mesh(pos, tri, normals=nrm, builder=DEVICE), rays)
    monkeypatch.delenv("MINIPATH_B200_DEVICE_LEAF")
    ctx.close()
    for k, v in out.items():
        print(f"QUALITY {name:7s} {k:28s} {v}")
    d, h = out["device (default leaf size)"], out["host reference split"]
    assert abs(d["hits"] - h["hits"]) <= 2e-3 * len(rays)  # same geometry up to the leaf-local quantisation
    assert d["sectors"] <= QUALITY_FACTOR * h["sectors"]
