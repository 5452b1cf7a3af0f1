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
