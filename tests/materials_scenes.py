"""Small scenes with a material table for the f3 tests (CPU and GPU sides share them)."""
import numpy as np


def _quad_grid(origin, eu, ev, nu, nv, uv_scale=1.0):
    """(nu x nv) quads spanning origin + u*eu + v*ev; uv = (u, v) * uv_scale.  Returns positions, uvs, triangles."""
    u, v = np.meshgrid(np.linspace(0.0, 1.0, nu + 1), np.linspace(0.0, 1.0, nv + 1))
    p = np.asarray(origin, np.float64) + u[..., None] * np.asarray(eu, np.float64) + v[..., None] * np.asarray(ev, np.float64)
    i = np.arange(nu)[None, :] + (nu + 1) * np.arange(nv)[:, None]
    a, b, c, d = i, i + 1, i + nu + 1, i + nu + 2
    t = np.stack([np.stack([a, b, d], -1), np.stack([a, d, c], -1)], 2).reshape(-1, 3)
    return p.reshape(-1, 3), (np.stack([u, v], -1) * uv_scale).reshape(-1, 2), t


def painted_room():
    """A checkered floor, a red back wall, a green tilted slab that shadows part of the floor and a blue smooth-shaded bump:
    four materials (ids 0..3), texture coordinates on every vertex, flat and interpolated normals.
    Returns dict(positions, normals, triangles, texcoords, tri_material, albedos, checker, checker_scale)."""
    P, N, T, UV, M = [], [], [], [], []
    base = 0

    def add(p, uv, t, mat, n=None):
        nonlocal base
        P.append(p), UV.append(uv), T.append(t + base), M.append(np.full(len(t), mat, np.uint32))
        N.append(np.zeros_like(p) if n is None else n)
        base += len(p)

    add(*_quad_grid((-4, 0, -4), (0, 0, 8), (8, 0, 0), 8, 8, uv_scale=2.0), mat=0)  # floor, normal +y (flat)
    add(*_quad_grid((-4, 0, -4), (8, 0, 0), (0, 5, 0), 6, 4), mat=1)  # back wall at z = -4, faces +z
    add(*_quad_grid((-1.5, 2.2, -1.0), (2.5, 0.4, 0), (0, 0.3, 2.0), 3, 3), mat=2)  # slab above the floor
    # smooth bump: a paraboloid cap with analytic normals
    u, v = np.meshgrid(np.linspace(-1, 1, 13), np.linspace(-1, 1, 13))
    h = 0.8 * np.clip(1.0 - (u * u + v * v), 0.0, None)
    p = np.stack([2.0 + 0.9 * u, 0.02 + h, 1.5 + 0.9 * v], -1).reshape(-1, 3)
    n = np.stack([1.6 * u / 0.9, np.ones_like(u), 1.6 * v / 0.9], -1).reshape(-1, 3)
    i = np.arange(12)[None, :] + 13 * np.arange(12)[:, None]
    t = np.stack([np.stack([i, i + 14, i + 1], -1), np.stack([i, i + 13, i + 14], -1)], 2).reshape(-1, 3)
    add(p, np.stack([u, v], -1).reshape(-1, 2) * 3.0, t, mat=3, n=n)
    return {
        "positions": np.concatenate(P).astype(np.float32), "normals": np.concatenate(N).astype(np.float32),
        "triangles": np.concatenate(T).astype(np.uint32), "texcoords": np.concatenate(UV).astype(np.float32),
        "tri_material": np.concatenate(M),
        "albedos": np.array([[0.9, 0.85, 0.7], [0.8, 0.15, 0.1], [0.2, 0.7, 0.25], [0.15, 0.3, 0.85]], np.float32),
        "checker": np.array([True, False, False, True]), "checker_scale": np.array([1.0, 1.0, 1.0, 2.5], np.float32),
    }


PAINTED_CAMERA = {"center": (0.5, 3.0, 7.5), "target": (0.0, 1.0, 0.0), "up": (0.0, 1.0, 0.0), "f_number": 8.0, "focus_distance": 8.0}


def single_floor(size=8.0, cells_per_unit=1.0):
    """One big square in the plane y = 0 (two triangles), uv = (x, z) + size/2 in world units, material 0 with the checker."""
    s = size / 2
    pos = np.array([[-s, 0, -s], [s, 0, -s], [s, 0, s], [-s, 0, s]], np.float32)
    uv = (pos[:, [0, 2]] + s).astype(np.float32)
    tri = np.array([[0, 2, 1], [0, 3, 2]], np.uint32)
    return pos, uv, tri
