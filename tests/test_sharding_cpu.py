"""CPU, world_size 2, gloo: the batch-shard / logits-gather logic of the N>1 path (the forward
itself has no collective)."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ivit_b200.sharding import gather_logits, shard_range


def test_shard_range_partitions():
    for total in (0, 1, 7, 256, 257):
        for world in (1, 2, 3, 8):
            parts = [shard_range(total, r, world) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == total
            assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in parts]
            assert max(sizes) - min(sizes) <= 1


def _per_image(x):                      # stand-in for the per-image forward: any batch-independent map
    return torch.stack([x.flatten(1).sum(1), x.flatten(1).max(1).values, x[:, 0, 0, 0] * 3.0], dim=1)


def _worker(rank, world, port, total, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(0)
        images = torch.randn(total, 3, 8, 8, generator=g)           # same global batch on every rank
        lo, hi = shard_range(total, rank, world)
        local = _per_image(images[lo:hi])
        full = gather_logits(local, total)
        ok = torch.equal(full, _per_image(images))
        q.put((rank, bool(ok), tuple(full.shape)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("total", [8, 7])
def test_gather_matches_single_rank_world2(total):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + total
    procs = [ctx.Process(target=_worker, args=(r, 2, port, total, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res), res
    assert all(shape == (total, 3) for _, _, shape in res)
