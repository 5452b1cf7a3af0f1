#!/usr/bin/env python
"""Triangle-soup traversal sweep (BASELINE.json configs[4]): incoherent rays against synthetic soups of
1e6 / 1e7 (/ 1e8) triangles, rays and hits resident in HBM.  One JSON line per (soup, kernel, layout):
Mrays/s, visits per ray, algorithmic bytes per ray and the fraction of the L2 / HBM roofs.

    python scripts/soup_sweep.py --sizes 1e6,1e7 --rays 4194304
--builders picks who makes the tree: `device` (mp_bvh_build_device, csrc/build_device.cu), `median` (host median split, builder=1) or
`reference` (the reference's agglomerative split on all host threads, builder=0; minutes above 1e6 triangles).  All three emit the
reference node format, so the traversal kernels and their parity tests are the same; the line names the builder.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1e6,1e7")
    ap.add_argument("--rays", type=int, default=1 << 22)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--builders", default="device,median")
    ap.add_argument("--kernels", default="thread-decoded,thread-quantised,pool-quantised,pool-decoded")
    args = ap.parse_args()
    import torch

    import minipath_b200 as mp
    from minipath_b200 import scenes

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    ctx = mp.Context(0)
    stream = torch.cuda.Stream()
    ctx.set_stream(stream.cuda_stream)
    l2 = ctx.probe_l2_gbs(32 << 20)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    builder_ids = {"reference": 0, "median": 1, "device": 2}
    builder_names = {"reference": "reference split on the host (builder=0)", "median": "median split on the host (builder=1)", "device": "device builder (mp_bvh_build_device)"}
    kernels = {"thread-decoded": ("thread", mp.TRACE_THREAD, mp.LAYOUT_DECODED), "thread-quantised": ("thread", mp.TRACE_THREAD, mp.LAYOUT_QUANTISED),
               "pool-quantised": ("pool", mp.TRACE_POOL, mp.LAYOUT_QUANTISED), "pool-decoded": ("pool", mp.TRACE_POOL, mp.LAYOUT_DECODED)}
    for size, bname in [(int(float(x)), b) for x in args.sizes.split(",") for b in args.builders.split(",")]:
        pos, tri = scenes.soup_mesh(size)
        t0 = time.time()
        bvh = mp.TriangleBvh.from_mesh(pos, tri, builder=builder_ids[bname])
        build_s = time.time() - t0
        build_device_ms = getattr(bvh, "last_build_ms", None)
        st = bvh.stats()
        bmin, bmax = bvh.get_bounding_box()
        o, d = scenes.incoherent_rays(bmin, bmax, args.rays)
        d = d / np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
        with np.errstate(divide="ignore"):
            inv = np.where(d == 0, np.float32(np.inf), np.float32(1.0) / d).astype(np.float32)
        rays = torch.from_numpy(np.concatenate([o, d, inv], 1).astype(np.float32)).cuda()
        hits = torch.empty((args.rays, 4), dtype=torch.int32, device="cuda")
        scene_mb = (st["n_inner"] * 160 + st["n_packets"] * 192) / 1e6
        for vname, variant, layout in [kernels[k] for k in args.kernels.split(",")]:
            ctx.set_trace_variant(variant)
            ctx.set_node_layout(layout)
            ctx.upload_scene(bvh)
            with torch.cuda.stream(stream):
                ctx.enable_counters(True)
                ctx.reset_counters()
                ctx.trace_rays_device(rays.data_ptr(), args.rays, hits.data_ptr())
                stream.synchronize()
                cnt = ctx.counters()
                ctx.enable_counters(False)
                ctx.trace_rays_device(rays.data_ptr(), args.rays, hits.data_ptr())  # warm-up
                ms = []
                for _ in range(args.reps):
                    flush.fill_(1)
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(stream)
                    ctx.trace_rays_device(rays.data_ptr(), args.rays, hits.data_ptr())
                    e1.record(stream)
                    stream.synchronize()
                    ms.append(e0.elapsed_time(e1))
            t = min(ms) * 1e-3
            n = args.rays
            bytes_ray = (cnt["inner_visits"] * 160 + cnt["packet_visits"] * 144 + cnt["leaf_visits"] * 32) / n + 52
            hit_frac = float((hits[:, 0] != -1).float().mean())
            line = {"workload": f"soup-{size:.0e}", "triangles": size, "rays": n, "kernel": vname, "layout": "decoded" if layout == mp.LAYOUT_DECODED else "quantised",
                    "mrays_s": n / t / 1e6, "ms": t * 1e3, "inner_per_ray": cnt["inner_visits"] / n, "packets_per_ray": cnt["packet_visits"] / n,
                    "bytes_per_ray": bytes_ray, "hit_fraction": hit_frac, "scene_mb": scene_mb, "bvh": {k: st[k] for k in ("n_inner", "n_leaves", "n_packets", "max_depth")},
                    "algorithmic_gbs": bytes_ray * n / t / 1e9, "frac_of_l2": bytes_ray * n / t / 1e9 / l2, "frac_of_hbm": bytes_ray * n / t / 1e9 / hbm,
                    "l2_probe_gbs": l2, "hbm_gbs": hbm, "builder": builder_names[bname], "build_s": build_s, "build_device_ms": build_device_ms,
                    "leaves_per_ray": cnt["leaf_visits"] / n, "leaf_fill": st["mean_leaf_fill"], "l2": "flushed between reps"}
            print(json.dumps(line), flush=True)
        del rays, hits
    ctx.close()


if __name__ == "__main__":
    main()
