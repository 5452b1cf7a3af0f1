"""Host BVH builder (host_bvh_build.inc, following building.rs:83-207): the fork-join build used for large scenes must emit
exactly the arrays of the sequential recursion (parents before children, sibling subtrees in child order), for the
reference split (builder=0) and for the median split (builder=1)."""
import os

import numpy as np
import pytest

import minipath_b200 as mp
from minipath_b200 import scenes


def _arrays(bvh):
    return [np.ascontiguousarray(a).tobytes() for a in (bvh.inner_bounds(), bvh.inner_links(), bvh.inner_boxes(), bvh.packets(), bvh.packet_boxes(),
                                                        bvh.tri_shading(), bvh.vertices())]


def _build(pos, tri, builder, threads):
    old = os.environ.get("MINIPATH_B200_BUILD_THREADS")
    os.environ["MINIPATH_B200_BUILD_THREADS"] = str(threads)
    try:
        return mp.TriangleBvh.from_mesh(pos, tri, builder=builder)
    finally:
        if old is None:
            del os.environ["MINIPATH_B200_BUILD_THREADS"]
        else:
            os.environ["MINIPATH_B200_BUILD_THREADS"] = old


@pytest.mark.parametrize("n,builder", [(70_000, 1), (300_000, 1), (1_200_000, 1), (70_000, 0)])
def test_parallel_build_equals_sequential_build(n, builder):
    # 70k is just above the 65536-triangle threshold of the fork-join path; 1.2M also runs the threaded root split
    pos, tri = scenes.soup_mesh(n, seed=n + builder)
    seq = _build(pos, tri, builder, 1)
    par = _build(pos, tri, builder, 7)
    assert seq.stats() == par.stats()
    assert seq.raw.root == par.raw.root
    for a, b in zip(_arrays(seq), _arrays(par)):
        assert a == b


def test_links_are_in_range_and_cover_every_triangle():
    pos, tri = scenes.soup_mesh(200_000, seed=3)
    bvh = _build(pos, tri, 1, 5)
    st = bvh.stats()
    links = np.asarray(bvh.inner_links()).reshape(-1)
    live = links[links != 0xFFFFFFF8]
    count, index = live & 7, live >> 3
    assert (index[count == 0] < st["n_inner"]).all()
    assert (index[count > 0] + count[count > 0] <= st["n_packets"]).all()
    # every inner node except the root is referenced once; leaves tile the packet array exactly
    assert np.array_equal(np.sort(index[count == 0]), np.arange(1, st["n_inner"]))
    order = np.argsort(index[count > 0])
    starts, counts = index[count > 0][order], count[count > 0][order]
    assert starts[0] == 0 and np.array_equal(starts[1:], (starts + counts)[:-1]) and starts[-1] + counts[-1] == st["n_packets"]
    sh = np.asarray(bvh.tri_shading()).view(np.uint32).reshape(-1, 4)  # v0, v1, v2, flags|material; padding lanes are all zero
    real = sh[sh[:, 1] != 0]                                           # soup_mesh: triangle i owns vertices 3i..3i+2
    assert len(real) == 200_000 and len(np.unique(real[:, 0] // 3)) == 200_000


def test_stats_rejects_malformed_trees():
    """mp_bvh_stats is what mp_scene_upload / mp_render validate a caller-made MpFlatBvh with: links out of range, a node with two
    parents (a cycle), leaves that overlap and vertex ids out of range are MP_EINVAL, not a host crash."""
    import ctypes as C
    import os

    import minipath_b200 as mp
    from minipath_b200 import _ffi

    bvh = mp.TriangleBvh.with_obj(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "data", "teapot.obj"))
    good = bvh.stats()
    assert good["n_inner"] == bvh.raw.n_inner and good["n_packets"] == bvh.raw.n_packets
    links = bvh.inner_links()  # writable view of the library-owned array
    sh = bvh.tri_shading()

    def bad():
        s = _ffi.MpBvhStats()
        rc = _ffi.lib.mp_bvh_stats(bvh._ptr, C.byref(s))
        return rc

    lane = int(np.nonzero(links[0] != _ffi.LINK_NULL)[0][0])
    keep = int(links[0, lane])
    for wrong in ((bvh.raw.n_inner + 5) << 3, 0 << 3, ((bvh.raw.n_packets - 1) << 3) | 3):  # out of range, back to the root (cycle), leaf past the end
        links.flags.writeable = True
        links[0, lane] = wrong
        assert bad() == _ffi.MP_EINVAL, hex(wrong)
        links[0, lane] = keep
    assert bad() == _ffi.MP_OK
    other = [int(k) for k in links[0] if int(k) != _ffi.LINK_NULL and int(k) != keep]
    if other:
        links[0, lane] = other[0]  # two parents for one subtree
        assert bad() == _ffi.MP_EINVAL
        links[0, lane] = keep
    sh.flags.writeable = True
    v = int(sh[0, 0])
    sh[0, 0] = bvh.raw.n_vertices
    assert bad() == _ffi.MP_EINVAL
    sh[0, 0] = v
    assert bad() == _ffi.MP_OK
