"""mu-sharding across ranks: world_size-2 gloo processes on CPU."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from finitedifference_b200 import sharding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("count,world", [(0, 1), (1, 1), (9, 2), (9, 4), (24, 8), (7, 8), (256, 8), (5, 3)])
def test_shard_range_partitions(count, world):
    blocks = [sharding.shard_range(count, r, world) for r in range(world)]
    assert [i for b in blocks for i in b] == list(range(count))          # contiguous, ordered, complete
    sizes = [len(b) for b in blocks]
    assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)


def test_shard_range_rejects_bad_ranks():
    for args in [(4, 2, 2), (4, -1, 2), (4, 0, 0), (-1, 0, 1)]:
        with pytest.raises(ValueError):
            sharding.shard_range(*args)


def test_lattice_is_the_reference_training_set_for_nine():
    pts = sharding.mu_lattice(9)
    ref = [[a, b] for a in np.linspace(4.25, 5.5, 3) for b in np.linspace(0.015, 0.03, 3)]   # train_autoencoder.py:63-72
    assert np.allclose(pts, ref) and len(sharding.mu_lattice(48)) == 48
    assert all(4.25 <= a <= 5.5 and 0.015 <= b <= 0.03 for a, b in sharding.mu_lattice(200))


def _worker(rank, world, port, total, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mus = sharding.mu_lattice(total)
        mine = sharding.shard_mus(mus, rank, world)
        # the per-rank work is independent; the only exchanges are the timing reduction and,
        # optionally, gathering small per-sample results
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        t = sharding.max_over_ranks(10.0 + rank)
        units = torch.tensor([len(mine)], dtype=torch.int64)
        dist.all_reduce(units)
        if rank == 0:
            torch.save({"gathered": gathered, "t": t, "units": int(units.item()), "mus": mus}, out)
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
@pytest.mark.parametrize("total", [9, 48])
def test_two_ranks_cover_the_sweep_once(tmp_path, total):
    out = str(tmp_path / "r0.pt")
    port = 29500 + (os.getpid() % 2000) + total
    mp.spawn(_worker, args=(2, port, total, out), nprocs=2, join=True)
    res = torch.load(out)
    flat = [m for block in res["gathered"] for m in block]
    assert flat == res["mus"] and res["units"] == total
    assert res["t"] == 11.0                     # max over ranks, what bench.py reports
