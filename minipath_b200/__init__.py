"""minipath_b200 — host-side mirror of minipath's public API over the B200 renderer core.

Names follow the reference crate (src/lib.rs:8-10): ``render``, ``RenderSettings``, ``RenderProgress``, ``Camera``,
``Scene``, ``TriangleBvh``, ``ScreenBlock``.  Everything here is a thin layer over the C ABI in
``include/minipath_b200.h``; the device work happens in ``libminipath_b200.so`` (hand-written CUDA for sm_100a).
There is no CPU fallback: device calls raise ``MpError`` when no B200 is present.
"""
import ctypes as C
import math

import numpy as np

from . import _ffi
from ._ffi import (  # noqa: F401  (re-exported)
    LAYOUT_AUTO,
    LAYOUT_DECODED,
    LAYOUT_QUANTISED,
    MODE_AUTO,
    MODE_MEGAKERNEL,
    MODE_WAVEFRONT,
    NO_HIT,
    ORDER_NEAR_FIRST,
    ORDER_REFERENCE,
    SHARD_SPP,
    SHARD_TILES,
    TRACE_HYBRID,
    TRACE_HYBRID_POOL,
    TRACE_HYBRID_POOL_S0,
    TRACE_POOL,
    TRACE_OCTET,
    TRACE_PERSIST,
    TRACE_THREAD,
    MpError,
    check,
    lib,
)

from ._abi_types import HIT_DTYPE, RAY_DTYPE, TILE_DTYPE  # noqa: E402,F401


def _f3(v):
    return (C.c_float * 3)(float(v[0]), float(v[1]), float(v[2]))


# ---------------------------------------------------------------------------------------------------------
# node links (reference src/scene/triangle_bvh/mod.rs:57-114)
# ---------------------------------------------------------------------------------------------------------
class CompressedNodeLink:
    NULL = _ffi.LINK_NULL
    MAX_INDEX = _ffi.LINK_MAX_INDEX
    MIN_COUNT = 1
    MAX_COUNT = _ffi.LINK_MAX_COUNT

    @staticmethod
    def new_leaf(index, count):
        out = C.c_uint32()
        check(lib.mp_link_new_leaf(index, count, C.byref(out)))
        return out.value

    @staticmethod
    def new_inner(index):
        out = C.c_uint32()
        check(lib.mp_link_new_inner(index, C.byref(out)))
        return out.value

    @staticmethod
    def decode(link):
        kind, index, count = C.c_int(), C.c_uint32(), C.c_uint32()
        check(lib.mp_link_decode(link, C.byref(kind), C.byref(index), C.byref(count)))
        return ("null", "inner", "leaf")[kind.value], index.value, count.value


# ---------------------------------------------------------------------------------------------------------
# screen tiling (reference src/screen_block.rs)
# ---------------------------------------------------------------------------------------------------------
class ScreenBlock:
    """AABB<Point2<u32>>: half-open pixel rectangle [min, max)."""

    __slots__ = ("min", "max")

    def __init__(self, min_xy, max_xy):
        self.min = (int(min_xy[0]), int(min_xy[1]))
        self.max = (int(max_xy[0]), int(max_xy[1]))

    @classmethod
    def with_size(cls, origin, size):
        return cls(origin, (origin[0] + size[0], origin[1] + size[1]))

    def width(self):
        return self.max[0] - self.min[0]

    def height(self):
        return self.max[1] - self.min[1]

    def is_empty(self):  # screen_block.rs:10-12
        return not (self.min[0] < self.max[0] and self.min[1] < self.max[1])

    def area(self):
        return 0 if self.is_empty() else self.width() * self.height()

    def contains(self, p):
        return self.min[0] <= p[0] < self.max[0] and self.min[1] <= p[1] < self.max[1]

    def internal_points(self):
        """Iterator over the (x, y) pairs inside the block in C order, x fastest; exact length at every step (screen_block.rs:28-39,84-143)."""
        return InternalPoints(self)

    def tile_ordering(self, tile_size, seed=0):
        """Tiles of tile_size (clipped at the right/bottom), centre-out with Exp noise (screen_block.rs:46-81)."""
        arr = tile_ordering_array(self.min[0], self.min[1], self.max[0], self.max[1], tile_size, seed)
        return [ScreenBlock((t["min_x"], t["min_y"]), (t["max_x"], t["max_y"])) for t in arr]

    def __eq__(self, o):
        return isinstance(o, ScreenBlock) and self.min == o.min and self.max == o.max

    def __hash__(self):
        return hash((self.min, self.max))

    def __repr__(self):
        return f"ScreenBlock({self.min}, {self.max})"


class InternalPoints:
    """screen_block.rs:84-143: a cursor over the block, an ExactSizeIterator (``len()`` is exact before and during iteration)."""

    def __init__(self, block):
        if block.is_empty():  # InternalPoints::empty (:92-101)
            self.min_x, self.max, self.cursor = 1, (0, 0), [0, 0]
        else:
            self.min_x, self.max, self.cursor = block.min[0], block.max, [block.min[0], block.min[1]]

    def __iter__(self):
        return self

    def __next__(self):
        if self.cursor[1] >= self.max[1]:
            raise StopIteration
        ret = (self.cursor[0], self.cursor[1])
        self.cursor[0] += 1
        if self.cursor[0] >= self.max[0]:
            self.cursor[0] = self.min_x
            self.cursor[1] += 1
        return ret

    def __len__(self):
        if self.cursor[1] >= self.max[1]:
            return 0
        return (self.max[1] - self.cursor[1] - 1) * (self.max[0] - self.min_x) + (self.max[0] - self.cursor[0])


def tile_ordering_array(min_x, min_y, max_x, max_y, tile_size, seed=0):
    n = C.c_size_t()
    check(lib.mp_tile_ordering(min_x, min_y, max_x, max_y, tile_size, seed, None, 0, C.byref(n)))
    arr = np.zeros(n.value, dtype=TILE_DTYPE)
    if n.value:
        check(lib.mp_tile_ordering(min_x, min_y, max_x, max_y, tile_size, seed, arr.ctypes.data_as(C.POINTER(_ffi.MpTile)), n.value, C.byref(n)))
    return arr


# ---------------------------------------------------------------------------------------------------------
# camera (reference src/camera.rs)
# ---------------------------------------------------------------------------------------------------------
class Camera:
    """Thin-lens camera; every builder method returns a new Camera (camera.rs:54-121)."""

    def __init__(self, raw=None):
        self._c = _ffi.MpCamera()
        if raw is None:
            check(lib.mp_camera_default(C.byref(self._c)))
        else:
            C.memmove(C.byref(self._c), C.byref(raw), C.sizeof(_ffi.MpCamera))

    @classmethod
    def default(cls):
        return cls()

    def _copy(self):
        return Camera(self._c)

    def focus_distance(self, d):
        assert d >= 0.0
        c = self._copy()
        c._c.focus_distance = d
        return c

    def sensor_width(self, w):
        assert w > 0.0
        c = self._copy()
        c._c.sensor_size, c._c.sensor_is_width = w, 1
        return c

    def sensor_height(self, h):
        assert h > 0.0
        c = self._copy()
        c._c.sensor_size, c._c.sensor_is_width = h, 0
        return c

    def focal_length(self, f):
        c = self._copy()
        c._c.focal_length = f
        return c

    def f_number(self, n):
        assert n > 0.0
        c = self._copy()
        c._c.f_number = n
        return c

    def look_at(self, center, look_at, up):
        c = self._copy()
        check(lib.mp_camera_look_at(C.byref(c._c), _f3(center), _f3(look_at), _f3(up)))
        return c

    def look_direction(self, center, forward, up):
        c = self._copy()
        check(lib.mp_camera_look_direction(C.byref(c._c), _f3(center), _f3(forward), _f3(up)))
        return c

    def center_forward_up_right(self):
        return tuple(np.array(list(getattr(self._c, k)), dtype=np.float32) for k in ("center", "forward", "up", "right"))

    def build_sampler(self, resolution):
        s = _ffi.MpCameraSampler()
        check(lib.mp_camera_build_sampler(C.byref(self._c), int(resolution[0]), int(resolution[1]), C.byref(s)))
        return s


class RenderSettings:
    """renderer/mod.rs:7-13 plus extension fields whose defaults reproduce the reference (0 bounces, OS seed)."""

    def __init__(self, tile_size, sample_count, resolution, max_bounces=0, seed=0, mode=None, order=ORDER_REFERENCE,
                 tiles_per_batch=0, trace_kernel=None):
        if tile_size <= 0 or sample_count <= 0:
            raise ValueError("tile_size and sample_count are NonZeroU32")
        self.tile_size, self.sample_count = int(tile_size), int(sample_count)
        self.resolution = (int(resolution[0]), int(resolution[1]))
        self.max_bounces, self.seed, self.order = int(max_bounces), int(seed), int(order)
        self.mode = None if mode is None else int(mode)  # None: render() lets the library pick (MODE_AUTO); the per-context calls use the megakernel
        self.tiles_per_batch = int(tiles_per_batch)
        self.trace_kernel = trace_kernel  # None: chosen by the library; else TRACE_THREAD / TRACE_OCTET / TRACE_PERSIST

    def _raw(self):
        return _ffi.MpRenderSettings(self.tile_size, self.sample_count, self.resolution[0], self.resolution[1], self.max_bounces, MODE_AUTO if self.mode is None else self.mode,
                                     self.order, self.tiles_per_batch, 0 if self.trace_kernel is None else self.trace_kernel + 1, 0, self.seed)

    def params(self, sample_count=None, sample_base=0):
        return _ffi.MpRenderParams(self.sample_count if sample_count is None else sample_count, sample_base, self.max_bounces, MODE_MEGAKERNEL if self.mode is None else self.mode,
                                   self.order, self.tile_size, self.seed)


# ---------------------------------------------------------------------------------------------------------
# scene (reference src/scene/mod.rs, src/scene/triangle_bvh/)
# ---------------------------------------------------------------------------------------------------------
class TriangleBvh:
    """Compressed 8-wide BVH in the reference node format, flattened for the device (MpFlatBvh)."""

    def __init__(self, ptr):
        self._ptr = ptr

    @classmethod
    def with_obj(cls, path, quad_mode=0):
        """TriangleBvh::with_obj (building.rs:28-34). quad_mode=1 fan-triangulates polygons (deviation; cube.obj)."""
        out = C.POINTER(_ffi.MpFlatBvh)()
        check(lib.mp_bvh_build_obj(str(path).encode(), quad_mode, C.byref(out)))
        return cls(out)

    BUILDER_REFERENCE, BUILDER_MEDIAN, BUILDER_DEVICE = 0, 1, 2

    @classmethod
    def from_mesh(cls, positions, triangles, normals=None, builder=0, texcoords=None, tri_material=None, materials=None, device=0):
        """Indexed mesh → BVH.  ``builder``: 0 the reference's split on the host, 1 median split on the host, 2 the device builder
        (mp_bvh_build_device on GPU ``device``; ``last_build_ms`` then holds its device time).  ``texcoords`` (n_vertices, 2), ``tri_material`` (n_triangles,) ids into ``materials``, a
        MATERIAL_DTYPE array (see ``make_materials``); without ``materials`` the extension integrator shades a constant grey."""
        pos = np.ascontiguousarray(positions, dtype=np.float32).reshape(-1, 3)
        tri = np.ascontiguousarray(triangles, dtype=np.uint32).reshape(-1, 3)
        nrm = None if normals is None else np.ascontiguousarray(normals, dtype=np.float32).reshape(-1, 3)
        tex = None if texcoords is None else np.ascontiguousarray(texcoords, dtype=np.float32).reshape(-1, 2)
        tm = None if tri_material is None else np.ascontiguousarray(tri_material, dtype=np.uint32).reshape(-1)
        mats = None if materials is None else np.ascontiguousarray(materials, dtype=_ffi.MATERIAL_DTYPE).reshape(-1)
        if tex is not None and len(tex) != len(pos):
            raise ValueError("texcoords must have one uv pair per vertex")
        if tm is not None and len(tm) != len(tri):
            raise ValueError("tri_material must have one id per triangle")
        out = C.POINTER(_ffi.MpFlatBvh)()
        fp, up = C.POINTER(C.c_float), C.POINTER(C.c_uint32)
        if builder == cls.BUILDER_DEVICE:
            ms = C.c_float()
            check(lib.mp_bvh_build_device(device, pos.ctypes.data_as(fp), None if nrm is None else nrm.ctypes.data_as(fp),
                                          None if tex is None else tex.ctypes.data_as(fp), len(pos), tri.ctypes.data_as(up),
                                          None if tm is None else tm.ctypes.data_as(up), len(tri),
                                          None if mats is None else mats.ctypes.data, 0 if mats is None else len(mats), C.byref(out), C.byref(ms)))
            bvh = cls(out)
            bvh.last_build_ms = ms.value
            return bvh
        check(lib.mp_bvh_build_mesh_mat(pos.ctypes.data_as(fp), None if nrm is None else nrm.ctypes.data_as(fp),
                                        None if tex is None else tex.ctypes.data_as(fp), len(pos), tri.ctypes.data_as(up),
                                        None if tm is None else tm.ctypes.data_as(up), len(tri),
                                        None if mats is None else mats.ctypes.data, 0 if mats is None else len(mats), builder, C.byref(out)))
        return cls(out)

    def __del__(self):
        if getattr(self, "_ptr", None):
            lib.mp_bvh_free(self._ptr)
            self._ptr = None

    @property
    def raw(self):
        return self._ptr.contents

    def get_bounding_box(self):
        r = self.raw
        return np.array(list(r.bbox_min), np.float32), np.array(list(r.bbox_max), np.float32)

    def stats(self):
        s = _ffi.MpBvhStats()
        check(lib.mp_bvh_stats(self._ptr, C.byref(s)))
        return {k: getattr(s, k) for k, _ in s._fields_}

    def _view(self, field, count, dtype, shape):
        r = self.raw
        if count == 0:
            return np.zeros((0,) + shape, dtype)
        addr = C.cast(getattr(r, field), C.c_void_p).value
        buf = (C.c_char * (count * int(np.prod(shape)) * np.dtype(dtype).itemsize)).from_address(addr)
        return np.frombuffer(buf, dtype=dtype).reshape((count,) + shape)

    # numpy views of the flat arrays (borrowed; valid while self is alive)
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


def make_materials(albedos, checker=None, checker_scale=None):
    """MATERIAL_DTYPE table from (n, 3) albedos; ``checker`` marks the entries that get the procedural checker."""
    alb = np.asarray(albedos, np.float32).reshape(-1, 3)
    m = np.zeros(len(alb), _ffi.MATERIAL_DTYPE)
    m["albedo"] = alb
    m["checker_scale"] = _ffi.DEFAULT_CHECKER_SCALE if checker_scale is None else np.asarray(checker_scale, np.float32)
    if checker is not None:
        m["flags"] = np.where(np.asarray(checker, bool), _ffi.MAT_CHECKER, 0).astype(np.uint32)
    return m


class Scene:
    """scene/mod.rs:13-15"""

    def __init__(self, object):
        self.object = object


# ---------------------------------------------------------------------------------------------------------
# render() and RenderProgress (reference src/renderer/machinery.rs)
# ---------------------------------------------------------------------------------------------------------
class _ImageView(np.ndarray):
    """ndarray view of the library-owned image that keeps its RenderProgress (the owner of the memory) alive."""

    def __new__(cls, arr, owner):
        obj = np.asarray(arr).view(cls)
        obj._owner = owner
        return obj

    def __array_finalize__(self, obj):
        self._owner = getattr(obj, "_owner", None)


class RenderProgressSnapshot:
    def __init__(self, finished, total):
        self.finished, self.total = finished, total

    def percent(self):
        return 100.0 * self.finished / self.total


class RenderProgress:
    def __init__(self, handle, settings, keepalive):
        self._h, self._settings, self._keep = handle, settings, keepalive

    def progress(self):
        s = _ffi.MpProgressSnapshot()
        check(lib.mp_progress_snapshot(self._h, C.byref(s)))
        return RenderProgressSnapshot(s.finished, s.total)

    def is_finished(self):
        out = C.c_int()
        check(lib.mp_progress_is_finished(self._h, C.byref(out)))
        return bool(out.value)

    def elapsed(self):
        out = C.c_double()
        check(lib.mp_progress_elapsed_ms(self._h, C.byref(out)))
        return out.value / 1e3

    def abort(self):
        check(lib.mp_progress_abort(self._h))

    def wait(self):
        check(lib.mp_progress_wait(self._h))

    def image(self, copy=None):
        """RGBA8 image as (H, W, 4) uint8 (machinery.rs:175-177).  The reference hands out the image itself behind its lock; so does
        this once the render is finished: a read-only view of the library's (pinned) image that keeps this RenderProgress alive.
        While workers are still writing tiles (or with copy=True) a snapshot copied under the image lock is returned instead."""
        w, h = self._settings.resolution
        if copy is None:
            copy = not self.is_finished()
        if copy:
            img = np.empty((h, w, 4), np.uint8)
            check(lib.mp_progress_image(self._h, img.ctypes.data, img.nbytes))
            return img
        ptr, nbytes = C.c_void_p(), C.c_size_t()
        check(lib.mp_progress_image_ptr(self._h, C.byref(ptr), C.byref(nbytes)))
        buf = (C.c_uint8 * nbytes.value).from_address(ptr.value)
        view = np.frombuffer(buf, dtype=np.uint8).reshape(h, w, 4)
        view.flags.writeable = False
        self._views = getattr(self, "_views", 0) + 1
        return _ImageView(view, self)

    def __del__(self):
        if getattr(self, "_h", None):
            lib.mp_progress_free(self._h)
            self._h = None


def render(scene, camera, settings, started_tile_callback=None, finished_tile_callback=None, devices=None):
    """minipath::render (machinery.rs:20-123): returns immediately; tiles render on one worker thread per GPU."""
    def _started(_user, tile):
        started_tile_callback(ScreenBlock((tile.min_x, tile.min_y), (tile.max_x, tile.max_y)))

    def _finished(_user, tile, snap):
        finished_tile_callback(ScreenBlock((tile.min_x, tile.min_y), (tile.max_x, tile.max_y)), RenderProgressSnapshot(snap.finished, snap.total))

    # no callback given → NULL function pointer: a ctypes trampoline per tile costs a GIL round trip from the worker thread
    # (768 tiles x 2 callbacks were 6 ms of a 9 ms frame)
    cb0 = _ffi.STARTED_FN(_started) if started_tile_callback else _ffi.STARTED_FN()
    cb1 = _ffi.FINISHED_FN(_finished) if finished_tile_callback else _ffi.FINISHED_FN()
    raw = settings._raw()
    dev_arr, n_dev = None, 0
    if devices:
        dev_arr, n_dev = (C.c_int * len(devices))(*devices), len(devices)
    handle = C.c_void_p()
    check(lib.mp_render(scene.object._ptr, C.byref(camera._c), C.byref(raw), dev_arr, n_dev, cb0, cb1, None, C.byref(handle)))
    return RenderProgress(handle, settings, (cb0, cb1, scene, dev_arr))


def save_image(path, rgba):
    """Write an (H, W, 4) uint8 image as PNG (or PPM when the path ends in .ppm).  The reference has no image writer."""
    img = np.ascontiguousarray(rgba, dtype=np.uint8)
    h, w, c = img.shape
    if c != 4:
        raise ValueError("expected an RGBA image")
    fn = lib.mp_image_write_ppm if str(path).lower().endswith(".ppm") else lib.mp_image_write_png
    check(fn(str(path).encode(), img.ctypes.data, w, h))


# ---------------------------------------------------------------------------------------------------------
# device context (one per GPU): the seam under Worker::render_tile / Object::intersect
# ---------------------------------------------------------------------------------------------------------
def device_count():
    n = C.c_int()
    rc = lib.mp_device_count(C.byref(n))
    return n.value if rc == 0 else 0


def _tiles_arg(tiles):
    if isinstance(tiles, np.ndarray) and tiles.dtype == TILE_DTYPE:
        arr = np.ascontiguousarray(tiles)
    else:
        arr = np.array([(t.min[0], t.min[1], t.max[0], t.max[1]) for t in tiles], dtype=TILE_DTYPE)
    return arr, arr.ctypes.data_as(C.POINTER(_ffi.MpTile)), len(arr)


class Context:
    def __init__(self, device=0, _borrowed=None):
        self._h = C.c_void_p()
        self._owned = _borrowed is None
        if _borrowed is None:
            check(lib.mp_ctx_create(device, C.byref(self._h)))
        else:
            self._h = _borrowed  # a group member's context: the group owns it
        self.resolution = None
        self.trace_variant = TRACE_THREAD  # what mp_ctx_create starts with

    def close(self):
        if getattr(self, "_h", None):
            if self._owned:
                lib.mp_ctx_destroy(self._h)
            self._h = None

    def stream(self):
        out = C.c_void_p()
        check(lib.mp_ctx_stream(self._h, C.byref(out)))
        return out.value

    __del__ = close

    def set_stream(self, cuda_stream_ptr):
        check(lib.mp_ctx_set_stream(self._h, cuda_stream_ptr))

    def set_trace_variant(self, variant):
        check(lib.mp_ctx_set_trace_variant(self._h, variant))
        self.trace_variant = variant

    def set_node_layout(self, layout):
        """LAYOUT_AUTO (default), LAYOUT_DECODED (f32 child boxes / triangle edges written at upload) or LAYOUT_QUANTISED (reference records)."""
        check(lib.mp_ctx_set_node_layout(self._h, layout))

    def sync(self):
        check(lib.mp_ctx_sync(self._h))

    def enable_counters(self, on=True):
        check(lib.mp_ctx_enable_counters(self._h, int(on)))

    def counters(self):
        c = _ffi.MpCounters()
        check(lib.mp_counters(self._h, C.byref(c)))
        return {k: getattr(c, k) for k, _ in c._fields_}

    def reset_counters(self):
        check(lib.mp_counters_reset(self._h))

    def last_device_ms(self):
        ms = C.c_float()
        check(lib.mp_last_device_ms(self._h, C.byref(ms)))
        return ms.value

    def probe_fp32_tflops(self):
        out = C.c_float()
        check(lib.mp_probe_fp32_tflops(self._h, C.byref(out)))
        return out.value

    def probe_l2_gbs(self, nbytes=32 << 20):
        out = C.c_float()
        check(lib.mp_probe_l2_gbs(self._h, nbytes, C.byref(out)))
        return out.value

    def probe_l2_gbs_shape(self, nbytes, blocks_per_sm, vec_bytes):
        out = C.c_float()
        check(lib.mp_probe_l2_gbs_shape(self._h, nbytes, blocks_per_sm, vec_bytes, C.byref(out)))
        return out.value

    def probe_l1_gather(self, n_records, sectors_per_record=5):
        """10^9 32-byte sectors per second the L1 pipe delivers to an incoherent per-lane gather of records."""
        out = C.c_float()
        check(lib.mp_probe_l1_gather(self._h, n_records, sectors_per_record, C.byref(out)))
        return out.value

    def upload_scene(self, bvh):
        check(lib.mp_scene_upload(self._h, bvh._ptr))

    def set_camera(self, sampler):
        check(lib.mp_camera_set(self._h, C.byref(sampler)))

    def create_film(self, width, height):
        check(lib.mp_film_create(self._h, width, height))
        self.resolution = (width, height)

    def bind_film(self, device_ptr, width, height):
        check(lib.mp_film_bind_external(self._h, device_ptr, width, height))
        self.resolution = (width, height)

    def clear_film(self):
        check(lib.mp_film_clear(self._h))

    def render_tiles(self, tiles, params):
        arr, ptr, n = _tiles_arg(tiles)
        check(lib.mp_render_tiles(self._h, ptr, n, C.byref(params)))

    def resolve_rgba8(self, spp_total, out=None):
        w, h = self.resolution
        img = np.empty((h, w, 4), np.uint8) if out is None else out
        check(lib.mp_film_resolve_rgba8(self._h, spp_total, img.ctypes.data))
        return img

    def resolve_device(self, spp_total, device_ptr):
        check(lib.mp_film_resolve_device(self._h, spp_total, device_ptr))

    def resolve_tiles_rgba8(self, tiles, tile_size, spp_total):
        arr, ptr, n = _tiles_arg(tiles)
        out = np.zeros((n, tile_size, tile_size, 4), np.uint8)
        check(lib.mp_film_resolve_tiles_rgba8(self._h, ptr, n, tile_size, spp_total, out.ctypes.data))
        return out

    def read_film(self):
        w, h = self.resolution
        film = np.empty((h, w, 4), np.float32)
        check(lib.mp_film_read_f32(self._h, film.ctypes.data))
        return film

    def trace_rays(self, rays, order=ORDER_REFERENCE):
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        hits = np.empty(len(rays), dtype=HIT_DTYPE)
        check(lib.mp_trace_rays(self._h, rays.ctypes.data, len(rays), order, hits.ctypes.data))
        return hits

    def intersect_rays(self, rays, order=ORDER_REFERENCE):
        """Object::intersect with hit finalisation: HitRecord {t, point, normal, material, texture_coords} per ray."""
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        out = np.empty(len(rays), dtype=_ffi.HIT_RECORD_DTYPE)
        check(lib.mp_intersect_rays(self._h, rays.ctypes.data, len(rays), order, out.ctypes.data))
        return out

    def trace_rays_device(self, dev_rays, n, dev_hits, order=ORDER_REFERENCE):
        """Object::intersect on rays already resident in HBM (device pointers, e.g. torch tensors); asynchronous."""
        check(lib.mp_trace_rays_device(self._h, dev_rays, n, order, dev_hits))

    def trace_rays_device_noalloc(self, dev_rays, n, dev_hits, order=ORDER_REFERENCE):
        """mp_trace_rays_device_noalloc: the pool traversal on a static device scratch — no allocation in the call (include/minipath_b200.h)."""
        check(lib.mp_trace_rays_device_noalloc(self._h, dev_rays, n, order, dev_hits))

    def generate_rays(self, tiles, params):
        arr, ptr, n = _tiles_arg(tiles)
        count = C.c_uint64()
        total = int(sum(int(t["max_x"] - t["min_x"]) * int(t["max_y"] - t["min_y"]) for t in arr)) * params.sample_count
        rays = np.empty(total, dtype=RAY_DTYPE)
        check(lib.mp_generate_rays(self._h, ptr, n, C.byref(params), rays.ctypes.data, total, C.byref(count)))
        assert count.value == total
        return rays


class Group:
    """The GPUs of one box as the workers of render()'s tile scheduler (machinery.rs:51-116,205-209) — mp_group_*.

    ``Group(devices=[0, 1, ...])``: one process, a host thread per device while a frame renders.
    ``Group.for_rank(device, rank, world, nccl_id, queue_name)``: one process per GPU (torchrun); rank 0 makes ``nccl_id`` with
    ``Group.unique_id()`` and the caller ships it.  The films are combined by one NCCL sum-reduce to rank 0."""

    def __init__(self, devices=None, _handle=None):
        self._h = C.c_void_p()
        if _handle is not None:
            self._h = _handle
        else:
            arr, n = (None, 0) if not devices else ((C.c_int * len(devices))(*devices), len(devices))
            check(lib.mp_group_create(arr, n, C.byref(self._h)))
        self.resolution = None

    @staticmethod
    def unique_id():
        buf = (C.c_uint8 * _ffi.NCCL_ID_BYTES)()
        check(lib.mp_group_unique_id(buf))
        return bytes(buf)

    @classmethod
    def for_rank(cls, device, rank, world, nccl_id=None, queue_name=None):
        h = C.c_void_p()
        idbuf = None if nccl_id is None else (C.c_uint8 * _ffi.NCCL_ID_BYTES).from_buffer_copy(nccl_id)
        check(lib.mp_group_create_rank(device, rank, world, idbuf, None if queue_name is None else queue_name.encode(), C.byref(h)))
        return cls(_handle=h)

    def close(self):
        if getattr(self, "_h", None):
            lib.mp_group_destroy(self._h)
            self._h = None

    __del__ = close

    def size(self):
        a, b = C.c_int(), C.c_int()
        check(lib.mp_group_size(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def context(self, local_index=0):
        h = C.c_void_p()
        check(lib.mp_group_ctx(self._h, local_index, C.byref(h)))
        ctx = Context(_borrowed=h)
        ctx.resolution = self.resolution
        return ctx

    def upload_scene(self, bvh):
        check(lib.mp_group_scene_upload(self._h, bvh._ptr))

    def set_camera(self, sampler):
        check(lib.mp_group_camera_set(self._h, C.byref(sampler)))

    def set_trace_variant(self, variant):
        check(lib.mp_group_set_trace_variant(self._h, variant))

    def create_film(self, width, height):
        check(lib.mp_group_film_create(self._h, width, height))
        self.resolution = (width, height)

    clear_film = lambda self: self.create_film(*self.resolution)  # noqa: E731  (same size again = clear)

    def render(self, tiles, params, shard=SHARD_TILES, min_batch=0):
        arr, ptr, n = _tiles_arg(tiles)
        check(lib.mp_group_render(self._h, ptr, n, C.byref(params), shard, min_batch))

    def reduce_film(self):
        check(lib.mp_group_reduce_film(self._h))

    def resolve_rgba8(self, spp_total, out=None):
        w, h = self.resolution
        img = np.empty((h, w, 4), np.uint8) if out is None else out
        check(lib.mp_group_resolve_rgba8(self._h, spp_total, img.ctypes.data))
        return img

    def read_film(self, local_index=0):
        w, h = self.resolution
        film = np.empty((h, w, 4), np.float32)
        check(lib.mp_group_read_film(self._h, local_index, film.ctypes.data))
        return film

    def sync(self):
        check(lib.mp_group_sync(self._h))

    def last_ms(self):
        ms = C.c_float()
        check(lib.mp_group_last_ms(self._h, C.byref(ms)))
        return ms.value

    def member_stats(self, local_index=0):
        b, t, ms = C.c_uint64(), C.c_uint64(), C.c_float()
        check(lib.mp_group_member_stats(self._h, local_index, C.byref(b), C.byref(t), C.byref(ms)))
        return {"batches": b.value, "tiles": t.value, "ms": ms.value}


class TileQueue:
    """Cross-process tile counter in shared memory (the AtomicUsize of machinery.rs:43,205-209)."""

    def __init__(self, name, create):
        self._h = C.c_void_p()
        self._owner = bool(create)
        check(lib.mp_tileq_open(name.encode(), int(create), C.byref(self._h)))

    def reset(self):
        check(lib.mp_tileq_reset(self._h))

    def next(self, batch=1):
        out = C.c_uint64()
        check(lib.mp_tileq_next(self._h, batch, C.byref(out)))
        return out.value

    def peek(self):
        out = C.c_uint64()
        check(lib.mp_tileq_peek(self._h, C.byref(out)))
        return out.value

    def barrier(self, world, timeout_ms=60000):
        check(lib.mp_tileq_barrier(self._h, world, timeout_ms))

    def close(self):
        if getattr(self, "_h", None):
            lib.mp_tileq_close(self._h, int(self._owner))
            self._h = None

    __del__ = close


def full_screen_tiles(width, height, tile_size, seed=1):
    return tile_ordering_array(0, 0, width, height, tile_size, seed)
