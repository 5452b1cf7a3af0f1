"""Plain ctypes / numpy mirror of the structs and constants of include/minipath_b200.h — NO library is loaded here.

``_ffi.py`` (the binding of libminipath_b200.so) takes its types from here; so may any other binding of the same C structs
(the test-side CPU checker loads this file by path, so that the CPU arms of bench.py never map the product library).
"""
import ctypes as C

import numpy as np

MP_OK, MP_EINVAL, MP_ECUDA, MP_ENODEVICE, MP_EIO, MP_ESTATE, MP_ENOMEM = 0, -1, -2, -3, -4, -5, -6
LINK_NULL = 0xFFFFFFF8
LINK_MAX_INDEX = 0x1FFFFFFE
LINK_MAX_COUNT = 7
NO_HIT = 0xFFFFFFFF
TRI_FLAT = 0x80000000
TRI_MATERIAL_MASK = 0x7FFFFFFF
MAT_CHECKER = 1
DEFAULT_ALBEDO = 0.7
DEFAULT_CHECKER_SCALE = 4.0
MODE_MEGAKERNEL, MODE_WAVEFRONT = 0, 1
ORDER_REFERENCE, ORDER_NEAR_FIRST = 0, 1
MODE_AUTO = 2
TRACE_THREAD, TRACE_OCTET, TRACE_PERSIST, TRACE_HYBRID, TRACE_POOL, TRACE_HYBRID_POOL, TRACE_HYBRID_POOL_S0 = 0, 1, 2, 3, 4, 5, 6
LAYOUT_QUANTISED, LAYOUT_DECODED, LAYOUT_AUTO = 0, 1, 2
SHARD_TILES, SHARD_SPP = 0, 1
NCCL_ID_BYTES = 128


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


class MpMaterial(C.Structure):
    _fields_ = [("albedo", C.c_float * 3), ("flags", C.c_uint32), ("checker_scale", C.c_float), ("_pad", C.c_float * 3)]


class MpFlatBvh(C.Structure):
    _fields_ = [
        ("bbox_min", C.c_float * 3),
        ("bbox_max", C.c_float * 3),
        ("root", C.c_uint32),
        ("n_inner", C.c_uint32),
        ("n_packets", C.c_uint32),
        ("n_vertices", C.c_uint32),
        ("n_triangles", C.c_uint32),
        ("n_materials", C.c_uint32),
        ("inner", C.POINTER(MpInnerNode)),
        ("inner_box", C.POINTER(MpBox)),
        ("packets", C.POINTER(MpTrianglePacket)),
        ("packet_box", C.POINTER(MpBox)),
        ("tri_shading", C.POINTER(MpTriShading)),
        ("vertices", C.POINTER(MpVertexShading)),
        ("materials", C.POINTER(MpMaterial)),
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


class MpHitRecord(C.Structure):
    _fields_ = [
        ("t", C.c_float),
        ("point", C.c_float * 3),
        ("normal", C.c_float * 3),
        ("material", C.c_uint32),
        ("texture_coords", C.c_float * 3),
        ("hit", C.c_int32),
    ]


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


RAY_DTYPE = np.dtype([("origin", "<f4", 3), ("direction", "<f4", 3), ("inv_direction", "<f4", 3)])
HIT_DTYPE = np.dtype([("tri", "<u4"), ("t", "<f4"), ("u", "<f4"), ("v", "<f4")])
HIT_RECORD_DTYPE = np.dtype([("t", "<f4"), ("point", "<f4", 3), ("normal", "<f4", 3), ("material", "<u4"), ("texture_coords", "<f4", 3), ("hit", "<i4")])
MATERIAL_DTYPE = np.dtype([("albedo", "<f4", 3), ("flags", "<u4"), ("checker_scale", "<f4"), ("_pad", "<f4", 3)])
TILE_DTYPE = np.dtype([("min_x", "<u4"), ("min_y", "<u4"), ("max_x", "<u4"), ("max_y", "<u4")])
