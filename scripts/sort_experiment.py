#!/usr/bin/env python
"""Would sorting bounce rays pay?  Bounce-like rays on the Sponza stand-in (origins on surfaces, directions uniform on the sphere)
traced by the pool kernel in three orders: random (what depth >= 2 of the wavefront sees), sorted by the Morton code of the origin,
and sorted by direction octant then Morton code.  The hits are order-independent, so a device-side sort would keep parity.

    python scripts/sort_experiment.py --rays 4194304
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def morton3(q):
    def spread(x):
        x = x.astype(np.uint64) & 0x3FF
        x = (x | (x << 16)) & 0x30000FF
        x = (x | (x << 8)) & 0x300F00F
        x = (x | (x << 4)) & 0x30C30C3
        x = (x | (x << 2)) & 0x9249249
        return x
    return spread(q[:, 0]) | (spread(q[:, 1]) << 1) | (spread(q[:, 2]) << 2)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rays", type=int, default=1 << 22)
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    import torch

    import minipath_b200 as mp
    from minipath_b200 import scenes

    bvh, _ = scenes.atrium_scene()
    bmin, bmax = (np.asarray(v, dtype=np.float32) for v in bvh.get_bounding_box())
    ctx = mp.Context(0)
    stream = torch.cuda.Stream()
    ctx.set_stream(stream.cuda_stream)
    ctx.set_trace_variant(mp.TRACE_POOL)
    ctx.upload_scene(bvh)
    n = args.rays

    def pack(o, d):
        d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
        with np.errstate(divide="ignore"):
            inv = np.where(d == 0, np.float32(np.inf), np.float32(1.0) / d).astype(np.float32)
        return np.concatenate([o, d, inv], 1).astype(np.float32)

    o, d = scenes.incoherent_rays(bmin, bmax, n)
    first = pack(o, d)
    hits = ctx.trace_rays(first.view(mp.RAY_DTYPE).reshape(-1))
    ok = hits["tri"] != 0xFFFFFFFF
    t = hits["t"][ok]
    p = first[ok, 0:3] + (t[:, None] - 1e-3 * float(np.linalg.norm(bmax - bmin))) * first[ok, 3:6]
    rng = np.random.Generator(np.random.PCG64(7))
    rays = pack(p.astype(np.float32), rng.normal(size=p.shape).astype(np.float32))
    m = len(rays)
    q = np.clip(((rays[:, 0:3] - bmin) / (bmax - bmin) * 1024).astype(np.int64), 0, 1023)
    code = morton3(q)
    octant = ((rays[:, 3] < 0).astype(np.uint64) | ((rays[:, 4] < 0).astype(np.uint64) << 1) | ((rays[:, 5] < 0).astype(np.uint64) << 2))
    orders = {"random": np.arange(m), "morton(origin)": np.argsort(code, kind="stable"), "octant,morton": np.argsort((octant << 30) | code, kind="stable"),
              "morton>>15,octant": np.argsort(((code >> 15) << 3) | octant, kind="stable")}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    dev_hits = torch.empty((m, 4), dtype=torch.int32, device="cuda")
    ref = None
    for name, perm in orders.items():
        dev = torch.from_numpy(rays[perm]).cuda()
        with torch.cuda.stream(stream):
            ctx.trace_rays_device(dev.data_ptr(), m, dev_hits.data_ptr())
            ms = []
            for _ in range(args.reps):
                flush.fill_(1)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                ctx.trace_rays_device(dev.data_ptr(), m, dev_hits.data_ptr())
                e1.record(stream)
                stream.synchronize()
                ms.append(e0.elapsed_time(e1))
        back = np.empty((m, 4), dtype=np.int32)
        back[perm] = dev_hits.cpu().numpy()
        if ref is None:
            ref = back
        print(json.dumps({"order": name, "rays": m, "ms": min(ms), "mrays_s": m / min(ms) / 1e3, "hits_equal_to_random_order": bool(np.array_equal(ref, back))}), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
