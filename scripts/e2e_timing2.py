import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import minipath_b200 as mp
from minipath_b200 import _ffi
import ctypes as C
bvh = mp.TriangleBvh.with_obj("data/teapot.obj")
cam = mp.Camera.default().look_at((0, 2, 10), (0, 1.5, 0), (0, 1, 0)).f_number(4.8).focus_distance(10.0)
w, h = 2048, 1536
st = mp.RenderSettings(64, 10, (w, h), max_bounces=0, seed=7, mode=mp.MODE_MEGAKERNEL)
ctx = mp.Context(0)
tiles = mp.full_screen_tiles(w, h, 64)
ctx.upload_scene(bvh); ctx.set_camera(cam.build_sampler((w, h))); ctx.create_film(w, h)
pin = C.c_void_p(); _ffi.check(_ffi.lib.mp_host_alloc_pinned(w*h*4, C.byref(pin)))
img_pinned = np.frombuffer((C.c_uint8 * (w*h*4)).from_address(pin.value), dtype=np.uint8)
img_page = np.empty(w*h*4, np.uint8)
def T(label, fn):
    t0 = time.perf_counter(); fn(); t1 = time.perf_counter(); print("%-28s %.3f ms" % (label, (t1-t0)*1e3))
for it in range(3):
    print("--- iter", it)
    T("clear_film (async)", ctx.clear_film)
    T("sync", ctx.sync)
    T("render_tiles (async launch)", lambda: ctx.render_tiles(tiles, st.params()))
    T("sync (kernel)", ctx.sync)
    print("   device ms", ctx.last_device_ms())
    T("resolve_rgba8 -> pinned", lambda: ctx.resolve_rgba8(10, out=img_pinned.reshape(h, w, 4)))
    T("resolve_rgba8 -> pageable", lambda: ctx.resolve_rgba8(10, out=img_page.reshape(h, w, 4)))
    time.sleep(0.05)
import threading
def seq():
    T("[thread] clear_film", ctx.clear_film)
    T("[thread] render_tiles (launch)", lambda: ctx.render_tiles(tiles, st.params()))
    T("[thread] sync (kernel)", ctx.sync)
    T("[thread] resolve -> pinned", lambda: ctx.resolve_rgba8(10, out=img_pinned.reshape(h, w, 4)))
    T("[thread] sync idle", ctx.sync)
for it in range(3):
    th = threading.Thread(target=seq); th.start(); th.join()
# range API as mp_render uses it
from minipath_b200 import _ffi as F
arr = tiles
def seq2():
    T("[range] tiles_upload", lambda: F.check(F.lib.mp_tiles_upload(ctx._h, tiles.ctypes.data_as(C.POINTER(F.MpTile)), len(tiles), 64)))
    prm = st.params()
    T("[range] render_tile_range", lambda: F.check(F.lib.mp_render_tile_range(ctx._h, 0, len(tiles), C.byref(prm))))
    T("[range] resolve -> pinned", lambda: ctx.resolve_rgba8(10, out=img_pinned.reshape(h, w, 4)))
for it in range(3):
    th = threading.Thread(target=seq2); th.start(); th.join()
def seq3():
    T("[full] upload_scene", lambda: ctx.upload_scene(bvh))
    T("[full] create_film", lambda: ctx.create_film(w, h))
    T("[full] tiles_upload", lambda: F.check(F.lib.mp_tiles_upload(ctx._h, tiles.ctypes.data_as(C.POINTER(F.MpTile)), len(tiles), 64)))
    prm = st.params()
    T("[full] render_tile_range", lambda: F.check(F.lib.mp_render_tile_range(ctx._h, 0, len(tiles), C.byref(prm))))
    T("[full] resolve -> pinned", lambda: ctx.resolve_rgba8(10, out=img_pinned.reshape(h, w, 4)))
    T("[full] sync idle", ctx.sync)
for it in range(3):
    th = threading.Thread(target=seq3); th.start(); th.join()
