"""N>1 host logic on CPU: world_size-2 gloo processes shard a frame (by samples, by static tiles, by the shared-memory
tile queue) and combine the films with one sum-reduce to rank 0, exactly as bench.py / the GPU path do with NCCL.
The per-rank renderer plugged in here is the CPU oracle (tests may use it); the thing under test is the sharding and
reduce plumbing of minipath_b200.dist and the cross-process TileQueue (machinery.rs:43,205-209)."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as tmp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H, SPP, TS = 96, 72, 5, 32


def _scene():
    import minipath_b200 as mp

    bvh = mp.TriangleBvh.with_obj(os.path.join(ROOT, "data", "teapot.obj"))
    cam = mp.Camera.default().look_at((0, 2, 10), (0, 1.5, 0), (0, 1, 0)).f_number(4.8).focus_distance(10.0)
    return bvh, cam.build_sampler((W, H)), mp.full_screen_tiles(W, H, TS, seed=11)


def _oracle_renderer(film_np, bvh, sampler):
    import minipath_b200 as mp
    from oracle_lib import Oracle

    oracle = Oracle()

    def render_fn(tiles, base, count):
        st = mp.RenderSettings(TS, count, (W, H), seed=5)
        oracle.render_tiles(bvh, sampler, W, H, tiles, st.params(sample_base=base), film=film_np, threads=2)

    return render_fn


def _worker(rank, world, port, scheme, qname, out_path):
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import minipath_b200 as mp
    from minipath_b200 import dist as mpdist

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    bvh, sampler, tiles = _scene()
    film = torch.zeros((H, W, 4), dtype=torch.float32)
    queue = None
    if qname:
        if rank == 0:
            queue = mp.TileQueue(qname, create=True)
            queue.reset()
        dist.barrier()
        if rank != 0:
            queue = mp.TileQueue(qname, create=False)
    units = mpdist.render_sharded(_oracle_renderer(film.numpy(), bvh, sampler), film, tiles, SPP, rank, world, scheme, queue=queue, batch=2,
                                  reduce_fn=lambda t: dist.reduce(t, dst=0, op=dist.ReduceOp.SUM))
    total = torch.tensor([units], dtype=torch.int64)
    dist.all_reduce(total)
    if rank == 0:
        np.save(out_path, film.numpy())
        assert int(total) == len(tiles) * SPP
    dist.barrier()
    if queue is not None:
        queue.close()
    dist.destroy_process_group()


def _run(scheme, tmp_path, qname=None, world=2):
    port = 29500 + (os.getpid() % 2000)
    out = str(tmp_path / "film.npy")
    tmp.spawn(_worker, args=(world, port, scheme, qname, out), nprocs=world, join=True)
    return np.load(out)


def test_split_helpers():
    from minipath_b200 import dist as mpdist

    for spp in (1, 5, 64, 1024):
        for world in (1, 2, 3, 8):
            parts = [mpdist.split_samples(spp, r, world) for r in range(world)]
            assert sum(c for _, c in parts) == spp
            assert parts[0][0] == 0 and all(parts[i][0] + parts[i][1] == parts[i + 1][0] for i in range(world - 1))
    for n in (0, 1, 7, 510):
        for world in (1, 2, 8):
            got = np.sort(np.concatenate([mpdist.split_tiles_static(n, r, world) for r in range(world)]))
            assert np.array_equal(got, np.arange(n))


def test_spp_split_reduce_two_ranks(tmp_path, oracle):
    """Sample split: film == film(samples of rank 0) + film(samples of rank 1), bit for bit."""
    import minipath_b200 as mp
    from minipath_b200 import dist as mpdist

    got = _run("spp", tmp_path)
    bvh, sampler, tiles = _scene()
    parts = []
    for r in range(2):
        base, count = mpdist.split_samples(SPP, r, 2)
        st = mp.RenderSettings(TS, count, (W, H), seed=5)
        f, _, _ = oracle.render_tiles(bvh, sampler, W, H, tiles, st.params(sample_base=base), threads=2)
        parts.append(f)
    assert np.array_equal((parts[0] + parts[1]).view(np.uint32), got.view(np.uint32))
    assert got[..., 3].sum() > 0


@pytest.mark.parametrize("dynamic", [False, True])
def test_tile_split_reduce_two_ranks(tmp_path, oracle, dynamic):
    """Tile split (static interleave, and the shared-memory dynamic queue): the reduced film equals the single-worker film."""
    import minipath_b200 as mp

    got = _run("tiles", tmp_path, qname=f"/mpq_test_{os.getpid()}" if dynamic else None)
    bvh, sampler, tiles = _scene()
    st = mp.RenderSettings(TS, SPP, (W, H), seed=5)
    ref, _, _ = oracle.render_tiles(bvh, sampler, W, H, tiles, st.params(), threads=2)
    assert np.array_equal(ref.view(np.uint32), got.view(np.uint32))
