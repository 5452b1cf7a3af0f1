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
