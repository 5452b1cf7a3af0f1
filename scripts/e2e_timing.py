import os, sys, time
sys.path.insert(0, "/root/repo")
import numpy as np
import minipath_b200 as mp
bvh = mp.TriangleBvh.with_obj("/root/repo/data/teapot.obj")
cam = mp.Camera.default().look_at((0, 2, 10), (0, 1.5, 0), (0, 1, 0)).f_number(4.8).focus_distance(10.0)
st = mp.RenderSettings(64, 10, (2048, 1536), max_bounces=0, seed=7, mode=mp.MODE_MEGAKERNEL)
for i in range(4):
    t0 = time.perf_counter(); prog = mp.render(mp.Scene(bvh), cam, st, devices=[0]); t1 = time.perf_counter(); prog.wait(); t2 = time.perf_counter(); img = prog.image(); t3 = time.perf_counter(); del prog, img; t4=time.perf_counter()
    print("render() %.3f wait %.3f image %.3f free %.3f ms" % ((t1-t0)*1e3, (t2-t1)*1e3, (t3-t2)*1e3, (t4-t3)*1e3), file=sys.stderr)
