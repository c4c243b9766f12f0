"""Splitting a parameter sweep over one process per GPU (first row).

The training-set sweep of the reference (``train_autoencoder.py:63-72``: a lattice over
mu1 in [4.25, 5.5] x mu2 in [0.015, 0.03], one ``load_or_compute_snaps`` call per point) is a set
of independent trajectories, so each rank takes a contiguous block of the list and no data-path
collective exists; ``torch.distributed`` is used for the rendez-vous and for reducing timings.
"""
from __future__ import annotations

import numpy as np

MU1_RANGE = (4.25, 5.5)      # config.py:25-27 of the reference
MU2_RANGE = (0.015, 0.03)


def shard_range(count: int, rank: int, world: int) -> range:
    """Contiguous block of ``range(count)`` owned by ``rank``: sizes differ by at most one, the
    first ``count % world`` ranks take the larger blocks, every index is owned exactly once."""
    if world < 1 or not 0 <= rank < world or count < 0:
        raise ValueError(f"bad shard request: count={count} rank={rank} world={world}")
    base, extra = divmod(count, world)
    lo = rank * base + min(rank, extra)
    return range(lo, lo + base + (1 if rank < extra else 0))


def mu_lattice(total: int, mu1_range=MU1_RANGE, mu2_range=MU2_RANGE):
    """``total`` parameter points on the smallest near-square lattice of the training range that
    holds them, mu1-major like the reference's nested loops; the 3x3 training set for total=9."""
    if total < 1:
        return []
    m1 = int(np.ceil(np.sqrt(total)))
    m2 = int(np.ceil(total / m1))
    pts = [[float(a), float(b)] for a in np.linspace(*mu1_range, m1) for b in np.linspace(*mu2_range, m2)]
    return pts[:total]


def shard_mus(mus, rank: int, world: int):
    """This rank's block of a parameter list."""
    r = shard_range(len(mus), rank, world)
    return [list(mus[i]) for i in r]


def max_over_ranks(value: float, device=None) -> float:
    """MAX-reduce a timing over the ranks (identity without an initialised process group)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device if dist.get_backend() == "nccl" else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
