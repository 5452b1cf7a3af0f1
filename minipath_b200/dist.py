"""Sharding the render over the ranks of one box (one process per GPU, torch.distributed).

The reference's only parallelism is one worker thread per core pulling tiles from an atomic counter
(src/renderer/machinery.rs:51-62,205-209).  Here the workers are GPUs in separate processes:

* ``scheme="spp"``   every rank renders all tiles with its own slice of the sample indices (distinct RNG counters,
                     ``sample_base``), the f32x4 films are summed by one reduce to rank 0  (BASELINE config 4).
* ``scheme="tiles"`` the tile list is split over ranks, either statically (interleaved) or dynamically through the
                     shared-memory ``TileQueue`` (the AtomicUsize of machinery.rs:43); each rank writes only its
                     tiles into a zeroed full-resolution film, so the same sum-reduce is a gather  (BASELINE config 3).

No collective runs during traversal; the single exchange step is the film reduce (NCCL over NVLink on the GPU box,
gloo in the CPU tests).  ``render_fn(tiles, sample_base, sample_count)`` is whatever accumulates into ``film`` —
``Context.render_tiles`` bound to a film tensor on the GPU; the CPU tests plug the oracle in to check the plumbing.
"""
import numpy as np


def split_samples(sample_count, rank, world):
    """Contiguous slice of sample indices for this rank: (sample_base, count). Remainder goes to the low ranks."""
    q, r = divmod(int(sample_count), int(world))
    count = q + (1 if rank < r else 0)
    base = rank * q + min(rank, r)
    return base, count


def split_tiles_static(n_tiles, rank, world):
    """Interleaved static split: rank r takes tiles r, r+world, … (tile_ordering is centre-out, so this balances)."""
    return np.arange(rank, n_tiles, world, dtype=np.int64)


def render_sharded(render_fn, film, tiles, sample_count, rank, world, scheme="spp", queue=None, batch=32, reduce_fn=None):
    """Render this rank's share into ``film`` and combine on rank 0.

    render_fn(tiles_subset, sample_base, sample_count) accumulates into ``film`` (already zeroed by the caller).
    reduce_fn(film) performs the sum-reduce to rank 0 (``torch.distributed.reduce``); None for world == 1.
    Returns the number of (tile, sample) units this rank rendered.
    """
    units = 0
    if scheme == "spp":
        base, count = split_samples(sample_count, rank, world)
        if count:
            render_fn(tiles, base, count)
            units = len(tiles) * count
    elif scheme == "tiles":
        if queue is None:
            mine = split_tiles_static(len(tiles), rank, world)
            if len(mine):
                render_fn(tiles[mine], 0, sample_count)
            units = len(mine) * sample_count
        else:
            while True:
                first = queue.next(batch)
                if first >= len(tiles):
                    break
                sub = tiles[first:first + batch]
                render_fn(sub, 0, sample_count)
                units += len(sub) * sample_count
    else:
        raise ValueError(f"unknown sharding scheme {scheme!r}")
    if world > 1 and reduce_fn is not None:
        reduce_fn(film)
    return units
