"""Format invariants of a BVH8 in the reference node format, checked against the source mesh — shared by the CPU test of the host
builders (tests/test_independent.py) and the GPU test of the device builder (tests/test_gpu_builder.py)."""
import numpy as np

NULL = 0xFFFFFFF8
K = np.float32(1.0) / np.float32(65535.0)


def _dequant(q, size, lo):
    """fma(size, float(q) * K, min) in f32 — evaluated in f64 and rounded once: exact product + sum fits f64 for these magnitudes."""
    rel = (q.astype(np.float32) * K).astype(np.float64)
    return (size.astype(np.float64) * rel + lo.astype(np.float64)).astype(np.float32)


def check_builder_invariants(pos, tri, bvh):
    """Invariants of the reference node format on `bvh` against the SOURCE mesh (positions, triangles)."""
    r = bvh.raw
    links, bounds, ibox = bvh.inner_links(), bvh.inner_bounds(), bvh.inner_boxes()
    packets, pbox, shading = bvh.packets(), bvh.packet_boxes(), bvh.tri_shading()
    n_tri = len(tri)
    assert r.n_triangles == n_tri

    # ---- every source triangle exactly once; padding lanes are all-zero records ----
    real = np.zeros(r.n_packets * 8, bool)
    stack, leaves, seen_inner = [(int(r.root), None)], [], set()
    root_box = np.concatenate([np.array(list(r.bbox_min), np.float32), np.array(list(r.bbox_max), np.float32) - np.array(list(r.bbox_min), np.float32)])
    boxes = {}  # link -> (min, size) the node must be decoded against
    boxes[int(r.root)] = (root_box[:3], root_box[3:])
    order = []
    while stack:
        link, parent = stack.pop()
        if link == NULL:
            continue
        if link & 7:
            leaves.append(link)
            continue
        i = link >> 3
        assert i not in seen_inner, "an inner node has two parents"
        seen_inner.add(i)
        if parent is not None:
            assert i > parent, "parents precede their children (building.rs:130-132)"
        order.append(i)
        kids = links[i]
        live = kids != NULL
        assert 1 <= live.sum() <= 8 and not np.any(live[np.argmin(live):]) if not live.all() else True, "children are packed from lane 0"
        bmin, bsize = boxes[link]
        assert np.array_equal(ibox[i][[0, 1, 2]], bmin) and np.array_equal(ibox[i][[4, 5, 6]], bsize), "inner_box table = the box the parent decodes for this child"
        assert not bounds[i][:, ~live].any(), "NULL lanes hold all-zero bounds (compressed_geometry.rs:41)"
        for lane in np.nonzero(live)[0]:
            cmin = _dequant(bounds[i][0:3, lane], bsize, bmin)
            cmax = _dequant(bounds[i][3:6, lane], bsize, bmin)
            assert np.all(cmin <= cmax)
            boxes[int(kids[lane])] = (cmin, (cmax - cmin).astype(np.float32))  # AABBSized: size = max - min (aabb.rs:292-303)
            stack.append((int(kids[lane]), i))
    assert len(seen_inner) == r.n_inner
    covered = np.zeros(r.n_packets, bool)
    tri_of_leaf = {}
    for link in leaves:
        first, count = link >> 3, link & 7
        assert 1 <= count <= 7 and not covered[first:first + count].any()
        covered[first:first + count] = True
        bmin, bsize = boxes[link]
        for p in range(first, first + count):
            assert np.array_equal(pbox[p][[0, 1, 2]], bmin) and np.array_equal(pbox[p][[4, 5, 6]], bsize), "packet_box table = the leaf's box"
        tri_of_leaf[link] = (first, count)
    assert covered.all(), "every packet belongs to exactly one leaf"

    # source triangles <-> (packet, lane): vertex ids are kept in tri_shading; a padding lane is {0,0,0,0} with zero geometry
    ids = shading[:, :3]
    geom_zero = ~packets.reshape(r.n_packets, 9, 8).transpose(0, 2, 1).reshape(-1, 9).any(axis=1)
    pad = geom_zero & ~shading.any(axis=1)
    # packets of a leaf are full except the last one: count real lanes per leaf
    n_real = 0
    keys = []
    for link, (first, count) in tri_of_leaf.items():
        lanes = np.arange(first * 8, (first + count) * 8)
        is_pad = pad[lanes]
        k = int((~is_pad).sum())
        if is_pad.any():
            assert not is_pad[:k].any() and is_pad[k:].all(), "padding only at the end of a leaf's last packet"
            assert k > (count - 1) * 8, "no packet is entirely padding"
        real[lanes[:k]] = True
        n_real += k
    assert n_real == n_tri, f"{n_real} triangle lanes for {n_tri} source triangles"
    got = np.sort(ids[real].astype(np.uint64) @ np.array([1, 1 << 21, 1 << 42], np.uint64))
    want = np.sort(tri.astype(np.uint64) @ np.array([1, 1 << 21, 1 << 42], np.uint64))
    assert np.array_equal(got, want), "the multiset of (v0, v1, v2) triples differs from the source: a triangle was dropped, duplicated or re-wound"

    # ---- every dequantised vertex within half a quantisation step of its source; every box on the way down encloses it ----
    q = packets.transpose(0, 3, 1, 2).reshape(-1, 3, 3)[real]                   # [tri][vertex][axis]
    bx = np.repeat(pbox, 8, axis=0)[real]
    lo, size = bx[:, None, 0:3], bx[:, None, 4:7]
    deq = _dequant(q, np.broadcast_to(size, q.shape), np.broadcast_to(lo, q.shape))
    src = pos[ids[real]]
    tol = 0.5 * size.astype(np.float64) / 65535.0 + 4e-7 * (np.abs(lo) + np.abs(size)).astype(np.float64) + 1e-30
    assert np.all(np.abs(deq.astype(np.float64) - src.astype(np.float64)) <= tol), "a vertex is further than half a quantisation step from its source"

    # round-out (building.rs:146-155; compressed_geometry.rs:122-131): the decompressed box of every node contains all SOURCE
    # vertices of the triangles below it — up to the one-ulp slack of fma(size, q*K, min) against (p - min) / size
    def subtree_bounds(link):
        if link & 7:
            first, count = tri_of_leaf[link]
            lanes = np.arange(first * 8, (first + count) * 8)
            lanes = lanes[real[lanes]]
            v = pos[ids[lanes]].reshape(-1, 3)
            return v.min(0), v.max(0)
        los, his = [], []
        for k in links[link >> 3]:
            if int(k) != NULL:
                a, b = subtree_bounds(int(k))
                los.append(a), his.append(b)
        lo_, hi_ = np.min(los, 0), np.max(his, 0)
        return lo_, hi_

    import sys
    sys.setrecursionlimit(10000)
    check = list(boxes.items()) if len(boxes) < 4000 else list(boxes.items())[:: max(1, len(boxes) // 4000)]
    for link, (bmin, bsize) in check:
        lo_, hi_ = subtree_bounds(link)
        slack = 2e-6 * (np.abs(bmin) + np.abs(bsize)) + 1e-30
        assert np.all(lo_ >= bmin - slack) and np.all(hi_ <= bmin + bsize + slack), f"node {link:#x}: decompressed box does not enclose its geometry"
