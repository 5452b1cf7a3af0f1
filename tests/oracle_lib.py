"""ctypes binding of oracle/liboracle.so — the CPU restatement used ONLY as the checker (tests, smoke, cpu_baseline)."""
import ctypes as C
import os

import numpy as np

from minipath_b200 import HIT_DTYPE, RAY_DTYPE, TILE_DTYPE, _ffi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_PATH = os.path.join(ROOT, "oracle", "liboracle.so")


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
        "orc_rng_init": (u32, [u32, u32, u64]),
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
    }
    for name, (res, args) in sig.items():
        fn = getattr(o, name)
        fn.restype, fn.argtypes = res, args
    return o


class Oracle:
    def __init__(self):
        self.o = load()

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
