"""Procedural inputs for the workloads whose geometry the reference does not ship.

* ``atrium_mesh`` — STAND-IN for Crytek Sponza (BASELINE.json configs[2..3]): the reference's ``data/Sponza/sponza.obj``
  is a missing large blob (``.MISSING_LARGE_BLOBS:1``) and the reference itself never renders it (``src/cli.rs:28``).
  About 258 k triangles of the same character: a long open-roofed hall, two storeys of colonnades with arches,
  gallery slabs, draped cloths and vases.  Every number quoted on it is labelled "stand-in".
* ``soup_mesh`` — random triangle soup of BASELINE.json configs[4].

Deterministic: numpy PCG64 streams with fixed seeds.  Host-side only (numpy); the BVH build is the library's.
"""
import numpy as np

ATRIUM_SEED = 0x5B0A2A


def _grid(nu, nv):
    """Triangle indices of an (nu+1) x (nv+1) vertex grid, row-major in v."""
    i = np.arange(nu)[None, :] + (nu + 1) * np.arange(nv)[:, None]
    a, b, c, d = i, i + 1, i + nu + 1, i + nu + 2
    return np.stack([np.stack([a, b, d], -1), np.stack([a, d, c], -1)], 2).reshape(-1, 3)


def _surface(fn, nu, nv):
    """Tessellate p(u,v) on [0,1]^2; normals from central differences."""
    u, v = np.meshgrid(np.linspace(0.0, 1.0, nu + 1), np.linspace(0.0, 1.0, nv + 1))
    p = fn(u, v)
    e = 1e-4
    n = np.cross(fn(u + e, v) - fn(u - e, v), fn(u, v + e) - fn(u, v - e))
    ln = np.linalg.norm(n, axis=-1, keepdims=True)
    n = np.where(ln > 0, n / np.maximum(ln, 1e-30), np.array([0.0, 1.0, 0.0]))
    return p.reshape(-1, 3), n.reshape(-1, 3), _grid(nu, nv), np.stack([u, v], -1).reshape(-1, 2)


class _Mesh:
    def __init__(self):
        self.p, self.n, self.t, self.uv, self.m, self.base, self.part = [], [], [], [], [], 0, 0

    def add(self, p, n, t, uv):
        self.p.append(p), self.n.append(n), self.t.append(t + self.base), self.uv.append(uv)
        self.m.append(np.full(len(t), self.part, np.uint32))
        self.base += len(p)

    def arrays(self):
        return (np.concatenate(self.p).astype(np.float32), np.concatenate(self.n).astype(np.float32),
                np.concatenate(self.t).astype(np.uint32))


# parts of the stand-in, in the order atrium_mesh() emits them: (name, albedo, checker cells per uv unit or 0)
ATRIUM_PARTS = (
    ("floor", (0.62, 0.58, 0.52), 24.0),
    ("walls", (0.78, 0.70, 0.58), 0.0),
    ("columns", (0.80, 0.78, 0.72), 0.0),
    ("arches", (0.72, 0.66, 0.60), 0.0),
    ("gallery", (0.55, 0.50, 0.46), 8.0),
    ("cloth_red", (0.75, 0.12, 0.10), 0.0),
    ("cloth_green", (0.14, 0.55, 0.18), 6.0),
    ("cloth_blue", (0.12, 0.20, 0.70), 0.0),
    ("vases", (0.30, 0.45, 0.55), 0.0),
)


def atrium_mesh():
    """Returns (positions[nv,3] f32, normals[nv,3] f32, triangles[nt,3] u32)."""
    return _atrium().arrays()


def atrium_mesh_materials():
    """The same stand-in with per-part materials: (positions, normals, triangles, texcoords[nv,2] f32, tri_material[nt] u32,
    materials as rows (albedo r, g, b, checker cells per uv unit or 0)) — SURVEY §8 f3."""
    m = _atrium()
    p, n, t = m.arrays()
    table = np.array([[*a, c] for _, a, c in ATRIUM_PARTS], np.float32)
    return p, n, t, np.concatenate(m.uv).astype(np.float32), np.concatenate(m.m), table


def _atrium():
    rng = np.random.Generator(np.random.PCG64(ATRIUM_SEED))
    m = _Mesh()
    L, W, H = 30.0, 12.0, 10.0  # x, z, y extents
    S = np.stack

    # floor with slightly uneven flagstones, long and short walls
    bump = rng.uniform(-0.01, 0.01, (65, 161))
    m.part = 0
    m.add(*_surface(lambda u, v: S([u * L - L / 2, 0.02 * np.sin(40 * u) * np.sin(36 * v) + 0 * u, v * W - W / 2], -1), 160, 64))
    m.p[-1][:, 1] += bump.reshape(-1)
    m.part = 1
    for z, s in ((-W / 2, 1.0), (W / 2, -1.0)):
        m.add(*_surface(lambda u, v, z=z, s=s: S([(u if s > 0 else 1 - u) * L - L / 2, v * H, z + 0.03 * s * np.sin(25 * u) * np.cos(9 * v)], -1), 160, 48))
    for x, s in ((-L / 2, 1.0), (L / 2, -1.0)):
        m.add(*_surface(lambda u, v, x=x, s=s: S([x + 0.03 * s * np.cos(14 * u) * np.sin(7 * v), v * H, (1 - u if s > 0 else u) * W - W / 2], -1), 64, 48))

    # two storeys of colonnades: fluted columns, arches between neighbours, gallery slabs
    xs = np.linspace(-L / 2 + 2.0, L / 2 - 2.0, 12)
    for level, (y0, y1) in enumerate(((0.0, 4.0), (4.6, 8.2))):
        for z in (-W / 2 + 2.2, W / 2 - 2.2):
            for x in xs:
                r = 0.32 - 0.04 * level
                m.part = 2
                m.add(*_surface(lambda u, v, x=x, z=z, r=r, y0=y0, y1=y1: S([
                    x + r * (1 + 0.06 * np.cos(16 * 2 * np.pi * u)) * (1 - 0.15 * v) * np.cos(2 * np.pi * u), y0 + v * (y1 - y0),
                    z - r * (1 + 0.06 * np.cos(16 * 2 * np.pi * u)) * (1 - 0.15 * v) * np.sin(2 * np.pi * u)], -1), 32, 24))
            for xa, xb in zip(xs[:-1], xs[1:]):
                cx, R = 0.5 * (xa + xb), 0.5 * (xb - xa)
                m.part = 3
                m.add(*_surface(lambda u, v, cx=cx, R=R, z=z, y1=y1: S([
                    cx - (R + 0.18 * np.cos(2 * np.pi * v)) * np.cos(np.pi * u), y1 + 0.45 * (R + 0.18 * np.cos(2 * np.pi * v)) * np.sin(np.pi * u),
                    z + 0.18 * np.sin(2 * np.pi * v)], -1), 16, 24))
    m.part = 4
    for z0, z1 in ((-W / 2, -W / 2 + 2.6), (W / 2 - 2.6, W / 2)):
        m.add(*_surface(lambda u, v, z0=z0, z1=z1: S([u * L - L / 2, 4.3 + 0 * u, z0 + v * (z1 - z0)], -1), 160, 8))
        m.add(*_surface(lambda u, v, z0=z0, z1=z1: S([(1 - u) * L - L / 2, 4.2 + 0 * u, z0 + v * (z1 - z0)], -1), 160, 8))

    # draped cloths hanging across the hall
    for k in range(6):
        x0 = -L / 2 + 3.0 + k * 4.6
        ph = rng.uniform(0, 2 * np.pi, 3)
        m.part = 5 + k % 3
        m.add(*_surface(lambda u, v, x0=x0, ph=ph: S([
            x0 + 1.6 * u + 0.15 * np.sin(9 * v + ph[0]), 8.6 - 3.2 * v - 0.9 * np.sin(np.pi * u) * (0.3 + v) + 0.08 * np.sin(14 * u + ph[1]),
            -2.4 + 4.8 * u * 0 + 2.4 * np.sin(np.pi * (u - 0.5)) + 0.2 * np.sin(11 * v + ph[2])], -1), 64, 64))

    # vases on the floor
    m.part = 8
    for k in range(16):
        cx, cz = -L / 2 + 1.5 + (k % 8) * 3.85, (-1.0 if k < 8 else 1.0) * 1.4
        m.add(*_surface(lambda u, v, cx=cx, cz=cz: S([
            cx + (0.25 + 0.22 * np.sin(np.pi * v) ** 2) * np.cos(2 * np.pi * u), 0.05 + 1.1 * v,
            cz - (0.25 + 0.22 * np.sin(np.pi * v) ** 2) * np.sin(2 * np.pi * u)], -1), 32, 32))
    return m


def atrium_camera_params(bbox_min, bbox_max):
    """camera defined relative to the bounding box so that it survives the stand-in (plain numbers, no library)."""
    lo, hi = np.asarray(bbox_min, np.float64), np.asarray(bbox_max, np.float64)
    c, s = 0.5 * (lo + hi), hi - lo
    long_axis = 0 if s[0] >= s[2] else 2
    el, ey = np.zeros(3), np.array([0.0, 1.0, 0.0])
    el[long_axis] = 1.0
    eye = c - 0.30 * s[long_axis] * el - 0.15 * s[1] * ey
    target = c + 0.30 * s[long_axis] * el - 0.10 * s[1] * ey
    return {"eye": tuple(eye), "target": tuple(target), "up": (0.0, 1.0, 0.0), "focus": float(np.linalg.norm(target - eye)),
            "sensor_width": 0.036, "focal_length": 0.024, "f_number": 16.0}


def atrium_camera(bbox_min, bbox_max):
    from . import Camera

    p = atrium_camera_params(bbox_min, bbox_max)
    return (Camera.default().sensor_width(p["sensor_width"]).focal_length(p["focal_length"]).f_number(p["f_number"]).focus_distance(p["focus"])
            .look_at(p["eye"], p["target"], p["up"]))


_ATRIUM_CACHE = {}


def atrium_scene(builder=0, materials=False, device=0):
    """(TriangleBvh, Camera) of the Sponza stand-in; the BVH uses the reference split unless builder=1.  ``materials``: with the
    per-part material table (same geometry, same tree)."""
    from . import TriangleBvh, make_materials

    key = (builder, bool(materials), device if builder == 2 else 0)
    if key not in _ATRIUM_CACHE:
        if materials:
            p, n, t, uv, tm, table = atrium_mesh_materials()
            mats = make_materials(table[:, :3], checker=table[:, 3] > 0, checker_scale=np.where(table[:, 3] > 0, table[:, 3], 1.0))
            bvh = TriangleBvh.from_mesh(p, t, normals=n, builder=builder, texcoords=uv, tri_material=tm, materials=mats, device=device)
        else:
            p, n, t = atrium_mesh()
            bvh = TriangleBvh.from_mesh(p, t, normals=n, builder=builder, device=device)
        _ATRIUM_CACHE[key] = (bvh, atrium_camera(*bvh.get_bounding_box()))
    return _ATRIUM_CACHE[key]


def soup_mesh(n_triangles, seed=None):
    """centres ~U([-1,1]^3), two edge vectors ~U([-l,l]^3), l = N^(-1/3)."""
    n = int(n_triangles)
    rng = np.random.Generator(np.random.PCG64(0xC0FFEE + int(round(np.log10(n))) if seed is None else seed))
    ell = n ** (-1.0 / 3.0)
    c = rng.uniform(-1.0, 1.0, (n, 3)).astype(np.float32)
    e1 = rng.uniform(-ell, ell, (n, 3)).astype(np.float32)
    e2 = rng.uniform(-ell, ell, (n, 3)).astype(np.float32)
    pos = np.stack([c, c + e1, c + e2], 1).reshape(-1, 3)
    tri = np.arange(3 * n, dtype=np.uint32).reshape(-1, 3)
    return pos, tri


def incoherent_rays(bbox_min, bbox_max, n, seed=0xBEEF):
    """origin ~U(bbox), direction uniform on the sphere (unnormalised inputs; Ray::new normalises)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    o = rng.uniform(np.asarray(bbox_min), np.asarray(bbox_max), (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    return o, d
