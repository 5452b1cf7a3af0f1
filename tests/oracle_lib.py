"""ctypes binding of oracle/liboracle.so — the CPU restatement used ONLY as the checker (tests, smoke, cpu_baseline)."""
import ctypes as C
import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_PATH = os.path.join(ROOT, "oracle", "liboracle.so")


def _load_by_path(name, path):
    """Import one file of the product package WITHOUT importing the package (whose __init__ maps libminipath_b200.so): the CPU arms
    of bench.py and the oracle-only tests must be able to run with the checker alone."""
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


# struct mirrors of include/minipath_b200.h (plain ctypes, no library) and the procedural meshes (plain numpy)
# (registered under its canonical name: if the product package is imported later in the same process it reuses THESE classes, so a
# BVH built by the oracle can be handed to the product's entry points without a ctypes type mismatch)
_ffi = _load_by_path("minipath_b200._abi_types", os.path.join(ROOT, "minipath_b200", "_abi_types.py"))
scenes = sys.modules.get("minipath_b200.scenes") or _load_by_path("_mp_scenes", os.path.join(ROOT, "minipath_b200", "scenes.py"))
HIT_DTYPE, RAY_DTYPE, TILE_DTYPE = _ffi.HIT_DTYPE, _ffi.RAY_DTYPE, _ffi.TILE_DTYPE


class OrcHitRecord(C.Structure):
    _fields_ = [
        ("t", C.c_float),
        ("point", C.c_float * 3),
        ("normal", C.c_float * 3),
        ("material", C.c_uint32),
        ("texture_coords", C.c_float * 3),
        ("hit", C.c_int),
    ]


def load():
    if not os.path.exists(ORACLE_PATH):
        raise ImportError(f"{ORACLE_PATH} missing: run `make` (or __graft_entry__.build())")
    o = C.CDLL(ORACLE_PATH)
    P, vp, f, u32, u64, i = C.POINTER, C.c_void_p, C.c_float, C.c_uint32, C.c_uint64, C.c_int
    sig = {
        "orc_compress_unit": (C.c_uint16, [f, i]),
        "orc_decompress_unit_rel": (f, [C.c_uint16]),
        "orc_decompress_coord": (f, [C.c_uint16, f, f]),
        "orc_ray_new": (None, [P(f), P(f), P(_ffi.MpRay)]),
        "orc_aabb8_intersect": (None, [vp, vp, P(_ffi.MpRay), f, vp, vp]),
        "orc_triangle8_intersect": (u32, [vp, P(_ffi.MpRay), vp, vp, vp]),
        "orc_fma_dot": (f, [P(f), P(f)]),
        "orc_fma_cross": (None, [P(f), P(f), P(f)]),
        "orc_bit_iter": (i, [u64, P(i)]),
        "orc_link_new_leaf": (i, [u32, u32, P(u32)]),
        "orc_link_new_inner": (i, [u32, P(u32)]),
        "orc_link_decode": (i, [u32, P(i), P(u32), P(u32)]),
        "orc_color_channel_to_u8": (C.c_uint8, [f]),
        "orc_pcg_hash": (u32, [u32]),
        "orc_rng_init": (None, [u32, u32, u64, P(u32)]),
        "orc_rng_next": (u32, [P(u32)]),
        "orc_set_simd": (None, [i]),
        "orc_get_simd": (i, []),
        "orc_trace_rays": (None, [P(_ffi.MpFlatBvh), vp, u64, u32, vp, P(_ffi.MpCounters), i]),
        "orc_intersect": (None, [P(_ffi.MpFlatBvh), P(_ffi.MpRay), u32, P(OrcHitRecord)]),
        "orc_sample_ray": (None, [P(_ffi.MpCameraSampler), u32, u32, P(u32), P(_ffi.MpRay)]),
        "orc_generate_rays": (u64, [P(_ffi.MpCameraSampler), u32, P(_ffi.MpTile), u32, P(_ffi.MpRenderParams), vp, u64]),
        "orc_render_tiles": (C.c_double, [P(_ffi.MpFlatBvh), P(_ffi.MpCameraSampler), u32, u32, P(_ffi.MpTile), u32, P(_ffi.MpRenderParams), vp,
                                         P(_ffi.MpCounters), i]),
        "orc_resolve_rgba8": (None, [vp, u32, u32, u32, vp]),
        "orc_bvh_build_obj": (i, [C.c_char_p, i, P(P(_ffi.MpFlatBvh))]),
        "orc_bvh_build_mesh": (i, [vp, vp, u32, vp, u64, P(P(_ffi.MpFlatBvh))]),
        "orc_bvh_build_mesh_mat": (i, [vp, vp, vp, u32, vp, vp, u64, vp, u32, P(P(_ffi.MpFlatBvh))]),
        "orc_intersect_rays": (None, [P(_ffi.MpFlatBvh), vp, u64, u32, vp, i]),
        "orc_bvh_free": (None, [P(_ffi.MpFlatBvh)]),
        "orc_build_error": (C.c_char_p, []),
        "orc_brute_force": (None, [P(_ffi.MpFlatBvh), vp, u64, vp, i]),
        "orc_camera_sampler": (None, [P(f), P(f), P(f), f, i, f, f, f, u32, u32, P(_ffi.MpCameraSampler)]),
        "orc_tiles": (u32, [u32, u32, u32, vp, u32]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(o, name)
        fn.restype, fn.argtypes = res, args
    return o


def materials_table(albedos, checker=None, checker_scale=None):
    """MATERIAL_DTYPE rows for Oracle.build_mesh (the test side's own constructor of the table)."""
    alb = np.asarray(albedos, np.float32).reshape(-1, 3)
    m = np.zeros(len(alb), _ffi.MATERIAL_DTYPE)
    m["albedo"] = alb
    m["checker_scale"] = 4.0 if checker_scale is None else np.asarray(checker_scale, np.float32)
    if checker is not None:
        m["flags"] = np.asarray(checker, bool).astype(np.uint32)
    return m


class OracleBvh:
    """A flat BVH made by the ORACLE's own loader and builder (oracle/oracle_build.inc) — a scene source that is not the code under
    test.  Duck-types minipath_b200.TriangleBvh where the tests need it (``_ptr``, ``raw``, array views, bounding box)."""

    def __init__(self, o, ptr):
        self._o, self._ptr = o, ptr

    def __del__(self):
        if getattr(self, "_ptr", None):
            self._o.orc_bvh_free(self._ptr)
            self._ptr = None

    @property
    def raw(self):
        return self._ptr.contents

    def get_bounding_box(self):
        r = self.raw
        return np.array(list(r.bbox_min), np.float32), np.array(list(r.bbox_max), np.float32)

    def _view(self, field, count, dtype, shape):
        if count == 0:
            return np.zeros((0,) + shape, dtype)
        addr = C.cast(getattr(self.raw, field), C.c_void_p).value
        buf = (C.c_char * (count * int(np.prod(shape)) * np.dtype(dtype).itemsize)).from_address(addr)
        return np.frombuffer(buf, dtype=dtype).reshape((count,) + shape)

    def inner_bounds(self):
        return self._view("inner", self.raw.n_inner, np.uint16, (64,))[:, :48].reshape(-1, 6, 8)

    def inner_links(self):
        return self._view("inner", self.raw.n_inner, np.uint32, (32,))[:, 24:]

    def inner_boxes(self):
        return self._view("inner_box", self.raw.n_inner, np.float32, (8,))

    def packets(self):
        return self._view("packets", self.raw.n_packets, np.uint16, (3, 3, 8))

    def packet_boxes(self):
        return self._view("packet_box", self.raw.n_packets, np.float32, (8,))

    def tri_shading(self):
        return self._view("tri_shading", self.raw.n_packets * 8, np.uint32, (4,))

    def vertices(self):
        return self._view("vertices", self.raw.n_vertices, np.float32, (8,))

    def materials(self):
        r = self.raw
        if r.n_materials == 0:
            return np.zeros(0, _ffi.MATERIAL_DTYPE)
        buf = (C.c_char * (r.n_materials * _ffi.MATERIAL_DTYPE.itemsize)).from_address(C.cast(r.materials, C.c_void_p).value)
        return np.frombuffer(buf, dtype=_ffi.MATERIAL_DTYPE)

    def stats(self):
        """Structural numbers of the tree, walked here in Python (the product's mp_bvh_stats is not used on this side)."""
        links, n_inner = self.inner_links(), self.raw.n_inner
        out = {"n_inner": int(n_inner), "n_packets": int(self.raw.n_packets), "n_triangles": int(self.raw.n_triangles), "n_leaves": 0, "max_depth": 0, "max_stack": 1}
        stack = [(int(self.raw.root), 1, 1)]  # link, depth, stack entries alive when it is popped
        while stack:
            link, depth, live = stack.pop()
            if link == _ffi.LINK_NULL:
                continue
            out["max_depth"] = max(out["max_depth"], depth)
            if link & 7:
                out["n_leaves"] += 1
                continue
            kids = [int(k) for k in links[link >> 3] if int(k) != _ffi.LINK_NULL]
            out["max_stack"] = max(out["max_stack"], live - 1 + len(kids))
            for j, k in enumerate(kids):  # child j is popped when j of its siblings are still below it
                stack.append((k, depth + 1, live - 1 + j + 1))
        return out


class Oracle:
    def __init__(self):
        self.o = load()

    # ---- scene sources that do not touch the product ----
    def build_obj(self, path, quad_mode=0):
        out = C.POINTER(_ffi.MpFlatBvh)()
        if self.o.orc_bvh_build_obj(str(path).encode(), quad_mode, C.byref(out)) != 0:
            raise RuntimeError("oracle builder: " + self.o.orc_build_error().decode())
        return OracleBvh(self.o, out)

    def build_mesh(self, positions, triangles, normals=None, texcoords=None, tri_material=None, materials=None):
        pos = np.ascontiguousarray(positions, np.float32).reshape(-1, 3)
        tri = np.ascontiguousarray(triangles, np.uint32).reshape(-1, 3)
        nrm = None if normals is None else np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
        tex = None if texcoords is None else np.ascontiguousarray(texcoords, np.float32).reshape(-1, 2)
        tm = None if tri_material is None else np.ascontiguousarray(tri_material, np.uint32).reshape(-1)
        mats = None if materials is None else np.ascontiguousarray(materials, _ffi.MATERIAL_DTYPE).reshape(-1)
        assert tex is None or len(tex) == len(pos)
        assert tm is None or len(tm) == len(tri)
        out = C.POINTER(_ffi.MpFlatBvh)()
        ptr = lambda a: None if a is None else a.ctypes.data  # noqa: E731
        if self.o.orc_bvh_build_mesh_mat(pos.ctypes.data, ptr(nrm), ptr(tex), len(pos), tri.ctypes.data, ptr(tm), len(tri), ptr(mats),
                                         0 if mats is None else len(mats), C.byref(out)) != 0:
            raise RuntimeError("oracle builder: " + self.o.orc_build_error().decode())
        return OracleBvh(self.o, out)

    def intersect_rays(self, bvh, rays, order=0, threads=8):
        """Object::intersect with finalisation for a batch: HIT_RECORD_DTYPE records."""
        rays = np.ascontiguousarray(rays, RAY_DTYPE)
        out = np.empty(len(rays), _ffi.HIT_RECORD_DTYPE)
        self.o.orc_intersect_rays(bvh._ptr, rays.ctypes.data, len(rays), order, out.ctypes.data, threads)
        return out

    def brute_force(self, bvh, rays, threads=8):
        rays = np.ascontiguousarray(rays, RAY_DTYPE)
        hits = np.empty(len(rays), HIT_DTYPE)
        self.o.orc_brute_force(bvh._ptr, rays.ctypes.data, len(rays), hits.ctypes.data, threads)
        return hits

    def camera_sampler(self, center, target, up, resolution, sensor_size=0.024, sensor_is_width=False, focal_length=0.05, f_number=9.0, focus_distance=float("inf")):
        """Camera::default().look_at(...).build_sampler(resolution) on the oracle side (camera.rs:41-52,93-146)."""
        f3 = lambda v: (C.c_float * 3)(*[float(x) for x in v])  # noqa: E731
        s = _ffi.MpCameraSampler()
        self.o.orc_camera_sampler(f3(center), f3(target), f3(up), sensor_size, int(sensor_is_width), focal_length, f_number, focus_distance,
                                  int(resolution[0]), int(resolution[1]), C.byref(s))
        return s

    def tiles(self, width, height, tile_size):
        n = self.o.orc_tiles(width, height, tile_size, None, 0)
        arr = np.zeros(n, TILE_DTYPE)
        self.o.orc_tiles(width, height, tile_size, arr.ctypes.data, n)
        return arr

    def ray_new(self, origin, direction):
        r = _ffi.MpRay()
        self.o.orc_ray_new((C.c_float * 3)(*origin), (C.c_float * 3)(*direction), C.byref(r))
        return r

    def rays_from(self, origins, directions):
        out = np.zeros(len(origins), RAY_DTYPE)
        for k, (o, d) in enumerate(zip(origins, directions)):
            r = self.ray_new(o, d)
            out[k] = (tuple(r.origin), tuple(r.direction), tuple(r.inv_direction))
        return out

    def aabb8(self, bmin, bmax, ray, max_t):
        bmin = np.ascontiguousarray(bmin, np.float32).reshape(3, 8)
        bmax = np.ascontiguousarray(bmax, np.float32).reshape(3, 8)
        t1, t2 = np.zeros(8, np.float32), np.zeros(8, np.float32)
        self.o.orc_aabb8_intersect(bmin.ctypes.data, bmax.ctypes.data, C.byref(ray), max_t, t1.ctypes.data, t2.ctypes.data)
        return t1, t2

    def triangle8(self, v, ray):
        v = np.ascontiguousarray(v, np.float32).reshape(3, 3, 8)
        t, u, w = (np.zeros(8, np.float32) for _ in range(3))
        mask = self.o.orc_triangle8_intersect(v.ctypes.data, C.byref(ray), t.ctypes.data, u.ctypes.data, w.ctypes.data)
        return mask, t, u, w

    def trace_rays(self, bvh, rays, order=0, threads=8, counters=None):
        rays = np.ascontiguousarray(rays, RAY_DTYPE)
        hits = np.empty(len(rays), HIT_DTYPE)
        c = _ffi.MpCounters() if counters is None else counters
        self.o.orc_trace_rays(bvh._ptr, rays.ctypes.data, len(rays), order, hits.ctypes.data, C.byref(c), threads)
        return hits

    def intersect(self, bvh, ray, order=0):
        rec = OrcHitRecord()
        self.o.orc_intersect(bvh._ptr, C.byref(ray), order, C.byref(rec))
        return rec

    def generate_rays(self, sampler, width, tiles, params):
        tiles = np.ascontiguousarray(tiles, TILE_DTYPE)
        n = int(sum(int(t["max_x"] - t["min_x"]) * int(t["max_y"] - t["min_y"]) for t in tiles)) * params.sample_count
        rays = np.empty(n, RAY_DTYPE)
        got = self.o.orc_generate_rays(C.byref(sampler), width, tiles.ctypes.data_as(C.POINTER(_ffi.MpTile)), len(tiles), C.byref(params),
                                       rays.ctypes.data, n)
        assert got == n
        return rays

    def render_tiles(self, bvh, sampler, width, height, tiles, params, film=None, threads=8, counters=None):
        tiles = np.ascontiguousarray(tiles, TILE_DTYPE)
        if film is None:
            film = np.zeros((height, width, 4), np.float32)
        c = _ffi.MpCounters() if counters is None else counters
        secs = self.o.orc_render_tiles(bvh._ptr, C.byref(sampler), width, height, tiles.ctypes.data_as(C.POINTER(_ffi.MpTile)), len(tiles),
                                       C.byref(params), film.ctypes.data, C.byref(c), threads)
        return film, secs, c

    def resolve_rgba8(self, film, spp_total):
        h, w, _ = film.shape
        out = np.empty((h, w, 4), np.uint8)
        self.o.orc_resolve_rgba8(np.ascontiguousarray(film, np.float32).ctypes.data, w, h, spp_total, out.ctypes.data)
        return out
