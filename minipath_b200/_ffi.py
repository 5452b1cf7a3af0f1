"""ctypes binding of include/minipath_b200.h (the C ABI of libminipath_b200.so).

The library is the product; this module only declares its structs and prototypes.  It raises at import time if
the shared library has not been built (``make`` / ``__graft_entry__.build()``) — there is no Python or CPU
fallback for the device path.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MINIPATH_B200_LIB") or os.path.join(_HERE, "libminipath_b200.so")  # env override: A/B builds of the same ABI

from ._abi_types import *  # noqa: F401,F403  (constants and struct mirrors, no library loaded there)
from ._abi_types import MpCamera, MpCameraSampler, MpCounters, MpFlatBvh, MpBvhStats, MpProgressSnapshot, MpRenderParams, MpRenderSettings, MpTile  # noqa: F401


class MpError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"minipath_b200 error {code}: {message}")
        self.code = code


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
    "mp_bvh_build_mesh_mat": (_i, [_P(C.c_float), _P(C.c_float), _P(C.c_float), _u32, _P(_u32), _P(_u32), _u64, _vp, _u32, _i, _P(_P(MpFlatBvh))]),
    "mp_bvh_build_device": (_i, [_i, _P(C.c_float), _P(C.c_float), _P(C.c_float), _u32, _P(_u32), _P(_u32), _u64, _vp, _u32, _P(_P(MpFlatBvh)), _P(C.c_float)]),
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
    "mp_trace_rays_device_noalloc": (_i, [_vp, _vp, _u64, _u32, _vp]),
    "mp_intersect_rays": (_i, [_vp, _vp, _u64, _u32, _vp]),
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
