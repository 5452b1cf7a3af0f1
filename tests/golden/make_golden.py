#!/usr/bin/env python
"""Writes tests/golden/teapot_gate.npz — a REGRESSION fixture of the CPU oracle, not a reference output.

The reference is a Rust crate that cannot be built in this image (no cargo / rustc), so there is no way to generate vectors from
it; what pins the oracle to the reference are the reference's own known-answer tests ported in tests/test_oracle_kat.py.  This file
freezes the oracle's answers on a small seeded case so that (a) a later change to oracle/ that alters a single bit is caught on
the CPU, and (b) the CUDA path is also checked against bytes committed to the repo, not only against an oracle built the same day.

    python tests/golden/make_golden.py        # run from the repo root; needs `make` to have built oracle/liboracle.so
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

W, H, TILE = 96, 72, 32


def case(mp):
    bvh = mp.TriangleBvh.with_obj(os.path.join(ROOT, "data", "teapot.obj"))
    cam = mp.Camera.default().look_at((0.0, 2.0, 10.0), (0.0, 1.5, 0.0), (0.0, 1.0, 0.0)).f_number(4.8).focus_distance(10.0)
    return bvh, cam.build_sampler((W, H)), mp.full_screen_tiles(W, H, TILE, seed=5)


def main():
    import minipath_b200 as mp
    from oracle_lib import Oracle

    oracle = Oracle()
    bvh, sampler, tiles = case(mp)
    st = mp.RenderSettings(TILE, 1, (W, H), max_bounces=0, seed=99)
    rays = oracle.generate_rays(sampler, W, tiles, st.params())
    hits = oracle.trace_rays(bvh, rays)
    out = {"rays": rays, "hits": hits}
    for bounces in (0, 5):
        stb = mp.RenderSettings(TILE, 4, (W, H), max_bounces=bounces, seed=99)
        film, _, _ = oracle.render_tiles(bvh, sampler, W, H, tiles, stb.params())
        out[f"film_b{bounces}"] = film
        out[f"film_b{bounces}_sha256"] = np.frombuffer(hashlib.sha256(film.tobytes()).digest(), np.uint8)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "teapot_gate.npz"), **out)
    print("wrote teapot_gate.npz:", len(rays), "rays,", int((hits["tri"] != mp.NO_HIT).sum()), "hits")


if __name__ == "__main__":
    main()
