"""The C ABI: libminipath_b200.so loads without a GPU and exports every function include/minipath_b200.h declares; the
ctypes binding declares exactly the same set; device entry points fail loudly (no CPU fallback) when no B200 is present."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_functions():
    text = open(os.path.join(ROOT, "include", "minipath_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mp_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from minipath_b200 import _ffi

    names = _header_functions()
    assert len(names) > 60
    lib = C.CDLL(_ffi.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in the header but not exported: {missing}"


def test_ctypes_binding_covers_the_header():
    from minipath_b200 import _ffi

    assert sorted(_ffi.PROTOTYPES) == _header_functions()


def test_no_cpu_fallback_without_device():
    import minipath_b200 as mp

    if mp.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(mp.MpError) as e:
        mp.Context(0)
    assert e.value.code == mp._ffi.MP_ENODEVICE
    bvh = mp.TriangleBvh.with_obj(os.path.join(ROOT, "data", "teapot.obj"))
    st = mp.RenderSettings(64, 1, (64, 64))
    with pytest.raises(mp.MpError):
        mp.render(mp.Scene(bvh), mp.Camera.default(), st)


def test_product_does_not_touch_the_oracle():
    """Nothing under minipath_b200/ may import, link or name the oracle (it is test infrastructure)."""
    pkg = os.path.join(ROOT, "minipath_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".inc")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "liboracle" not in text and "oracle_lib" not in text and "orc_" not in text, f


def test_render_rejects_bad_mode_and_kernel():
    """Invalid settings are MP_EINVAL (where the reference would panic), checked before any device is touched."""
    import minipath_b200 as mp

    bvh = mp.TriangleBvh.with_obj(os.path.join(ROOT, "data", "teapot.obj"))
    for kw in ({"mode": 3}, {"trace_kernel": mp.TRACE_HYBRID_POOL_S0 + 1}, {"order": 2}):
        with pytest.raises(mp.MpError) as e:
            mp.render(mp.Scene(bvh), mp.Camera.default(), mp.RenderSettings(64, 1, (64, 64), **kw))
        assert e.value.code == mp._ffi.MP_EINVAL, kw
    assert mp.RenderSettings(64, 1, (64, 64))._raw().mode == mp.MODE_AUTO
    assert mp.RenderSettings(64, 1, (64, 64)).params().mode == mp.MODE_MEGAKERNEL


def test_device_entry_points_fail_loudly_without_a_gpu():
    """No CPU fallback anywhere: on a box without a CUDA device the device entry points return an error code and a message."""
    import ctypes as C

    import numpy as np

    import minipath_b200 as mp
    from minipath_b200 import _ffi

    if mp.device_count() > 0:
        pytest.skip("this box has a GPU")
    lib = _ffi.lib
    out = C.c_void_p()
    assert lib.mp_ctx_create(0, C.byref(out)) == _ffi.MP_ENODEVICE and b"no CUDA device" in lib.mp_last_error()
    assert lib.mp_trace_rays_device_noalloc(None, None, 1, 0, None) == _ffi.MP_EINVAL
    pos = np.zeros((3, 3), np.float32)
    tri = np.array([[0, 1, 2]], np.uint32)
    bvh = C.POINTER(_ffi.MpFlatBvh)()
    fp, up = C.POINTER(C.c_float), C.POINTER(C.c_uint32)
    rc = lib.mp_bvh_build_device(0, pos.ctypes.data_as(fp), None, None, 3, tri.ctypes.data_as(up), None, 1, None, 0, C.byref(bvh), None)
    assert rc == _ffi.MP_ENODEVICE, lib.mp_last_error()
    dev = (C.c_int * 1)(0)
    grp = C.c_void_p()
    assert lib.mp_group_create(dev, 1, C.byref(grp)) != _ffi.MP_OK
    with pytest.raises(mp.MpError):
        mp.render(mp.Scene(mp.TriangleBvh.with_obj(os.path.join(ROOT, "data", "teapot.obj"))), mp.Camera.default(), mp.RenderSettings(64, 1, (64, 64))).wait()
