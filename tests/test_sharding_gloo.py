"""N > 1 host path on CPU: two `gloo` ranks shard a batch of independent frames with shard_slice(),
process their slices (the oracle stands in for the GPU kernels -- test infrastructure only), and the
gathered result equals the unsharded one.  Also checks the max-over-ranks timing reduction bench.py
uses and tile_rows()' row partition for the row-tiled path.  No GPU needed."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
import low_light_image_enhancement_b200 as L


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank, world, port, n_frames, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        frames = np.stack([oracle.synth_lowlight(24, 32, seed=100 + i) for i in range(n_frames)])
        sl = L.shard_slice(n_frames, rank, world)
        p = dict(sigmas=(2.0, 5.0), weights=(0.5, 0.5))
        mine = oracle.msrcr(frames[sl], p) if sl.stop > sl.start else np.zeros((0, 24, 32, 3), np.float32)
        # gather variable-length shards: pad to the largest shard
        counts = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(counts, torch.tensor([mine.shape[0]]))
        mx = int(max(c.item() for c in counts))
        pad = np.zeros((mx,) + mine.shape[1:], np.float32)
        pad[:mine.shape[0]] = mine
        bufs = [torch.zeros(pad.shape) for _ in range(world)]
        dist.all_gather(bufs, torch.from_numpy(pad))
        # bench.py's time reduction: max over ranks
        t = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            y = np.concatenate([b.numpy()[:int(c.item())] for b, c in zip(bufs, counts)])
            q.put((y, float(t.item()), [int(c.item()) for c in counts]))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_frames", [5, 2, 1])
def test_two_rank_gloo_sharding_matches_unsharded(n_frames):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_rank_main, args=(r, world, port, n_frames, q)) for r in range(world)]
    for p in procs:
        p.start()
    y, tmax, counts = q.get()
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    frames = np.stack([oracle.synth_lowlight(24, 32, seed=100 + i) for i in range(n_frames)])
    ref = oracle.msrcr(frames, dict(sigmas=(2.0, 5.0), weights=(0.5, 0.5)))
    assert sum(counts) == n_frames and max(counts) - min(counts) <= 1
    assert np.array_equal(y, ref)
    assert tmax == float(world)


def test_shard_slice_partitions_exactly():
    for n in (0, 1, 7, 64, 65):
        for world in (1, 2, 4, 8):
            sl = [L.shard_slice(n, r, world) for r in range(world)]
            assert sl[0].start == 0 and sl[-1].stop == n
            assert all(a.stop == b.start for a, b in zip(sl, sl[1:]))
            sizes = [s.stop - s.start for s in sl]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        L.shard_slice(4, 2, 2)


def test_tile_rows_partition():
    for h, parts in ((16384, 8), (1080, 4), (601, 3), (4096, 1), (777, 5)):
        t = L.tile_rows(h, parts)
        assert t[0][0] == 0 and sum(r for _, r in t) == h
        assert all(a[0] + a[1] == b[0] for a, b in zip(t, t[1:]))
        assert all(r0 % 4 == 0 for r0, _ in t)
    with pytest.raises(ValueError):
        L.tile_rows(8, 4)
    assert L.tile_rows(16384, 8) == [(2048 * d, 2048) for d in range(8)]
