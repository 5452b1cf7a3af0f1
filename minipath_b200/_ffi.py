"""ctypes binding of include/minipath_b200.h (the C ABI of libminipath_b200.so).

The library is the product; this module only declares its structs and prototypes.  It raises at import time if
the shared library has not been built (``make`` / ``__graft_entry__.build()``) — there is no Python or CPU
fallback for the device path.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MINIPATH_B200_LIB") or os.path.join(_HERE, "libminipath_b200.so")  # env override: A/B builds of the same ABI

MP_OK, MP_EINVAL, MP_ECUDA, MP_ENODEVICE, MP_EIO, MP_ESTATE, MP_ENOMEM = 0, -1, -2, -3, -4, -5, -6
LINK_NULL = 0xFFFFFFF8
LINK_MAX_INDEX = 0x1FFFFFFE
LINK_MAX_COUNT = 7
NO_HIT = 0xFFFFFFFF
TRI_FLAT = 0x80000000
MODE_MEGAKERNEL, MODE_WAVEFRONT = 0, 1
ORDER_REFERENCE, ORDER_NEAR_FIRST = 0, 1
MODE_AUTO = 2
TRACE_THREAD, TRACE_OCTET, TRACE_PERSIST, TRACE_HYBRID, TRACE_POOL, TRACE_HYBRID_POOL, TRACE_HYBRID_POOL_S0 = 0, 1, 2, 3, 4, 5, 6
LAYOUT_QUANTISED, LAYOUT_DECODED, LAYOUT_AUTO = 0, 1, 2
SHARD_TILES, SHARD_SPP = 0, 1
NCCL_ID_BYTES = 128


class MpError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"minipath_b200 error {code}: {message}")
        self.code = code


class MpBox(C.Structure):
    _fields_ = [("min", C.c_float * 3), ("_pad0", C.c_float), ("size", C.c_float * 3), ("_pad1", C.c_float)]


class MpInnerNode(C.Structure):
    _fields_ = [("bounds", (C.c_uint16 * 8) * 6), ("links", C.c_uint32 * 8)]


class MpTrianglePacket(C.Structure):
    _fields_ = [("v", ((C.c_uint16 * 8) * 3) * 3)]


class MpTriShading(C.Structure):
    _fields_ = [("v", C.c_uint32 * 3), ("flags_material", C.c_uint32)]


class MpVertexShading(C.Structure):
    _fields_ = [("normal", C.c_float * 3), ("_pad0", C.c_float), ("tex", C.c_float * 3), ("_pad1", C.c_float)]


class MpFlatBvh(C.Structure):
    _fields_ = [
        ("bbox_min", C.c_float * 3),
        ("bbox_max", C.c_float * 3),
        ("root", C.c_uint32),
        ("n_inner", C.c_uint32),
        ("n_packets", C.c_uint32),
        ("n_vertices", C.c_uint32),
        ("n_triangles", C.c_uint32),
        ("_pad", C.c_uint32),
        ("inner", C.POINTER(MpInnerNode)),
        ("inner_box", C.POINTER(MpBox)),
        ("packets", C.POINTER(MpTrianglePacket)),
        ("packet_box", C.POINTER(MpBox)),
        ("tri_shading", C.POINTER(MpTriShading)),
        ("vertices", C.POINTER(MpVertexShading)),
    ]


class MpBvhStats(C.Structure):
    _fields_ = [
        ("n_inner", C.c_uint32),
        ("n_leaves", C.c_uint32),
        ("n_packets", C.c_uint32),
        ("n_triangles", C.c_uint32),
        ("max_depth", C.c_uint32),
        ("max_stack", C.c_uint32),
        ("mean_leaf_fill", C.c_float),
        ("mean_inner_fill", C.c_float),
    ]


class MpCamera(C.Structure):
    _fields_ = [
        ("center", C.c_float * 3),
        ("forward", C.c_float * 3),
        ("up", C.c_float * 3),
        ("right", C.c_float * 3),
        ("focus_distance", C.c_float),
        ("sensor_size", C.c_float),
        ("sensor_is_width", C.c_int32),
        ("focal_length", C.c_float),
        ("f_number", C.c_float),
    ]


class MpCameraSampler(C.Structure):
    _fields_ = [
        ("center", C.c_float * 3),
        ("up", C.c_float * 3),
        ("right", C.c_float * 3),
        ("film_origin_offset", C.c_float * 3),
        ("pixel_scale", C.c_float),
        ("lens_radius", C.c_float),
        ("lens_weight", C.c_float),
    ]


class MpTile(C.Structure):
    _fields_ = [("min_x", C.c_uint32), ("min_y", C.c_uint32), ("max_x", C.c_uint32), ("max_y", C.c_uint32)]


class MpRay(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("direction", C.c_float * 3), ("inv_direction", C.c_float * 3)]


class MpHit(C.Structure):
    _fields_ = [("tri", C.c_uint32), ("t", C.c_float), ("u", C.c_float), ("v", C.c_float)]


class MpRenderParams(C.Structure):
    _fields_ = [
        ("sample_count", C.c_uint32),
        ("sample_base", C.c_uint32),
        ("max_bounces", C.c_uint32),
        ("mode", C.c_uint32),
        ("order", C.c_uint32),
        ("tile_size", C.c_uint32),
        ("seed", C.c_uint64),
    ]


class MpCounters(C.Structure):
    _fields_ = [
        ("rays", C.c_uint64),
        ("camera_rays", C.c_uint64),
        ("shadow_rays", C.c_uint64),
        ("inner_visits", C.c_uint64),
        ("packet_visits", C.c_uint64),
        ("leaf_visits", C.c_uint64),
        ("launches", C.c_uint64),
    ]


class MpRenderSettings(C.Structure):
    _fields_ = [
        ("tile_size", C.c_uint32),
        ("sample_count", C.c_uint32),
        ("resolution_x", C.c_uint32),
        ("resolution_y", C.c_uint32),
        ("max_bounces", C.c_uint32),
        ("mode", C.c_uint32),
        ("order", C.c_uint32),
        ("tiles_per_batch", C.c_uint32),
        ("trace_kernel", C.c_uint32),
        ("_pad", C.c_uint32),
        ("seed", C.c_uint64),
    ]


class MpProgressSnapshot(C.Structure):
    _fields_ = [("finished", C.c_size_t), ("total", C.c_size_t)]


STARTED_FN = C.CFUNCTYPE(None, C.c_void_p, MpTile)
FINISHED_FN = C.CFUNCTYPE(None, C.c_void_p, MpTile, MpProgressSnapshot)

_P = C.POINTER
_vp = C.c_void_p
_u32, _u64, _i = C.c_uint32, C.c_uint64, C.c_int

# name -> (restype, argtypes); every name declared in include/minipath_b200.h
PROTOTYPES = {
    "mp_last_error": (C.c_char_p, []),
    "mp_link_new_leaf": (_i, [_u32, _u32, _P(_u32)]),
    "mp_link_new_inner": (_i, [_u32, _P(_u32)]),
    "mp_link_decode": (_i, [_u32, _P(_i), _P(_u32), _P(_u32)]),
    "mp_bvh_build_obj": (_i, [C.c_char_p, _i, _P(_P(MpFlatBvh))]),
    "mp_bvh_build_mesh": (_i, [_P(C.c_float), _P(C.c_float), _u32, _P(_u32), _u64, _P(_P(MpFlatBvh))]),
    "mp_bvh_build_mesh_ex": (_i, [_P(C.c_float), _P(C.c_float), _u32, _P(_u32), _u64, _i, _P(_P(MpFlatBvh))]),
    "mp_bvh_free": (_i, [_P(MpFlatBvh)]),
    "mp_bvh_stats": (_i, [_P(MpFlatBvh), _P(MpBvhStats)]),
    "mp_camera_default": (_i, [_P(MpCamera)]),
    "mp_camera_look_at": (_i, [_P(MpCamera), _P(C.c_float), _P(C.c_float), _P(C.c_float)]),
    "mp_camera_look_direction": (_i, [_P(MpCamera), _P(C.c_float), _P(C.c_float), _P(C.c_float)]),
    "mp_camera_build_sampler": (_i, [_P(MpCamera), _u32, _u32, _P(MpCameraSampler)]),
    "mp_tile_ordering": (_i, [_u32, _u32, _u32, _u32, _u32, _u64, _P(MpTile), C.c_size_t, _P(C.c_size_t)]),
    "mp_device_count": (_i, [_P(_i)]),
    "mp_ctx_create": (_i, [_i, _P(_vp)]),
    "mp_ctx_destroy": (_i, [_vp]),
    "mp_ctx_pool_acquire": (_i, [_i, _P(_vp)]),
    "mp_ctx_pool_release": (_i, [_vp]),
    "mp_ctx_pool_clear": (_i, []),
    "mp_tiles_upload": (_i, [_vp, _P(MpTile), _u32, _u32]),
    "mp_render_tile_range": (_i, [_vp, _u32, _u32, _P(MpRenderParams)]),
    "mp_film_resolve_tiles_begin": (_i, [_vp, _u32, _u32, _u32, _u32, _i]),
    "mp_film_resolve_tiles_end": (_i, [_vp, _i, _P(_vp)]),
    "mp_ctx_set_stream": (_i, [_vp, _vp]),
    "mp_ctx_set_trace_variant": (_i, [_vp, _u32]),
    "mp_ctx_set_node_layout": (_i, [_vp, _u32]),
    "mp_ctx_sync": (_i, [_vp]),
    "mp_ctx_stream": (_i, [_vp, _P(_vp)]),
    "mp_ctx_enable_counters": (_i, [_vp, _i]),
    "mp_counters": (_i, [_vp, _P(MpCounters)]),
    "mp_counters_reset": (_i, [_vp]),
    "mp_last_device_ms": (_i, [_vp, _P(C.c_float)]),
    "mp_probe_fp32_tflops": (_i, [_vp, _P(C.c_float)]),
    "mp_probe_l2_gbs": (_i, [_vp, C.c_size_t, _P(C.c_float)]),
    "mp_probe_l2_gbs_shape": (_i, [_vp, C.c_size_t, _u32, _u32, _P(C.c_float)]),
    "mp_probe_l1_gather": (_i, [_vp, _u32, _u32, _P(C.c_float)]),
    "mp_scene_upload": (_i, [_vp, _P(MpFlatBvh)]),
    "mp_camera_set": (_i, [_vp, _P(MpCameraSampler)]),
    "mp_film_create": (_i, [_vp, _u32, _u32]),
    "mp_film_bind_external": (_i, [_vp, _vp, _u32, _u32]),
    "mp_film_clear": (_i, [_vp]),
    "mp_film_device_ptr": (_i, [_vp, _P(_vp)]),
    "mp_render_tiles": (_i, [_vp, _P(MpTile), _u32, _P(MpRenderParams)]),
    "mp_film_resolve_rgba8": (_i, [_vp, _u32, _vp]),
    "mp_film_resolve_device": (_i, [_vp, _u32, _vp]),
    "mp_film_resolve_tiles_rgba8": (_i, [_vp, _P(MpTile), _u32, _u32, _u32, _vp]),
    "mp_film_read_f32": (_i, [_vp, _vp]),
    "mp_trace_rays": (_i, [_vp, _vp, _u64, _u32, _vp]),
    "mp_trace_rays_device": (_i, [_vp, _vp, _u64, _u32, _vp]),
    "mp_generate_rays": (_i, [_vp, _P(MpTile), _u32, _P(MpRenderParams), _vp, _u64, _P(_u64)]),
    "mp_dev_alloc": (_i, [_vp, C.c_size_t, _P(_vp)]),
    "mp_dev_free": (_i, [_vp, _vp]),
    "mp_dev_upload": (_i, [_vp, _vp, _vp, C.c_size_t]),
    "mp_dev_download": (_i, [_vp, _vp, _vp, C.c_size_t]),
    "mp_host_alloc_pinned": (_i, [C.c_size_t, _P(_vp)]),
    "mp_host_free_pinned": (_i, [_vp]),
    "mp_render": (_i, [_P(MpFlatBvh), _P(MpCamera), _P(MpRenderSettings), _P(_i), _i, STARTED_FN, FINISHED_FN, _vp, _P(_vp)]),
    "mp_progress_snapshot": (_i, [_vp, _P(MpProgressSnapshot)]),
    "mp_progress_is_finished": (_i, [_vp, _P(_i)]),
    "mp_progress_elapsed_ms": (_i, [_vp, _P(C.c_double)]),
    "mp_progress_abort": (_i, [_vp]),
    "mp_progress_wait": (_i, [_vp]),
    "mp_progress_image": (_i, [_vp, _vp, C.c_size_t]),
    "mp_progress_image_ptr": (_i, [_vp, _P(_vp), _P(C.c_size_t)]),
    "mp_image_write_png": (_i, [C.c_char_p, _vp, _u32, _u32]),
    "mp_image_write_ppm": (_i, [C.c_char_p, _vp, _u32, _u32]),
    "mp_progress_free": (_i, [_vp]),
    "mp_tileq_open": (_i, [C.c_char_p, _i, _P(_vp)]),
    "mp_tileq_reset": (_i, [_vp]),
    "mp_tileq_next": (_i, [_vp, _u64, _P(_u64)]),
    "mp_tileq_peek": (_i, [_vp, _P(_u64)]),
    "mp_tileq_barrier": (_i, [_vp, _u32, _u32]),
    "mp_tileq_close": (_i, [_vp, _i]),
    "mp_group_create": (_i, [_P(_i), _i, _P(_vp)]),
    "mp_group_unique_id": (_i, [_vp]),
    "mp_group_create_rank": (_i, [_i, _i, _i, _vp, C.c_char_p, _P(_vp)]),
    "mp_group_destroy": (_i, [_vp]),
    "mp_group_size": (_i, [_vp, _P(_i), _P(_i)]),
    "mp_group_ctx": (_i, [_vp, _i, _P(_vp)]),
    "mp_group_scene_upload": (_i, [_vp, _P(MpFlatBvh)]),
    "mp_group_camera_set": (_i, [_vp, _P(MpCameraSampler)]),
    "mp_group_set_trace_variant": (_i, [_vp, _u32]),
    "mp_group_film_create": (_i, [_vp, _u32, _u32]),
    "mp_group_render": (_i, [_vp, _P(MpTile), _u32, _P(MpRenderParams), _u32, _u32]),
    "mp_group_reduce_film": (_i, [_vp]),
    "mp_group_resolve_rgba8": (_i, [_vp, _u32, _vp]),
    "mp_group_read_film": (_i, [_vp, _i, _vp]),
    "mp_group_sync": (_i, [_vp]),
    "mp_group_last_ms": (_i, [_vp, _P(C.c_float)]),
    "mp_group_member_stats": (_i, [_vp, _i, _P(_u64), _P(_u64), _P(C.c_float)]),
}


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make` (or __graft_entry__.build()). "
            "minipath_b200 has no fallback path without its CUDA library."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError here means the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def check(rc):
    if rc != MP_OK:
        raise MpError(rc, lib.mp_last_error().decode("utf-8", "replace"))
