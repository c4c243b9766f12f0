"""LSPG PROM with the grid rows dealt to the ranks of a process group, on CPU:
gloo processes drive the product's kernels compiled against the SIMT emulator.  The sharded
trajectory must reproduce the reference's (golden vectors from the unmodified reference) with the
same Gauss-Newton iteration counts, for any number of ranks."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden", "lspg_n24.npz")
TOL = 1e-8


def _run(rank, world, mode="reference", steps=None):
    from finitedifference_b200.rom import ShardedLspgSolver
    from oracle import burgers_oracle as orc
    from tests import emu_driver as E
    g = np.load(GOLDEN)
    n, steps = int(g["n"]), int(g["steps"]) if steps is None else steps
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, n, n)
    solver = ShardedLspgSolver(gx, gy, float(g["dt"]), list(g["mu"]), g["basis"], rank=rank, world=world,
                               jacobian=mode, _backend=(E.load(), torch.device("cpu")))
    snaps_own, red, hist = solver.run(np.ones(2 * n * n), steps)
    return solver, solver.gather_snapshots(snaps_own), red.numpy().T, [len(h) for h in hist], g


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        solver, snaps, red, its, g = _run(rank, world)
        torch.save({"snaps": snaps, "red": red, "its": its, "rows": solver.row_range,
                    "state_cells": int(solver.state_cells.size)}, f"{out}.{rank}")
    finally:
        dist.destroy_process_group()


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


def test_one_rank_equals_reference_golden():
    solver, snaps, red, its, g = _run(0, 1)
    assert solver.state_cells.size == 24 * 24
    assert its == g["gn_its"].tolist()
    assert max(rel(snaps[:, i], g["snaps"][:, i]) for i in range(snaps.shape[1])) < TOL


def test_one_rank_exact_operator_vs_oracle():
    from oracle import burgers_oracle as orc
    solver, snaps, red, its, g = _run(0, 1, mode="exact", steps=4)
    n = int(g["n"])
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, n, n)
    ref, _, red_o, hist = orc.lspg_trajectory(gx, gy, np.ones(2 * n * n), float(g["dt"]), 4, list(g["mu"]), g["basis"],
                                              jacobian_mode="exact", return_info=True)
    assert its == [len(h) for h in hist]
    assert max(rel(red[:, i], red_o[:, i]) for i in range(5)) < TOL


@pytest.mark.timeout(300)
@pytest.mark.parametrize("world", [2, 3])
def test_ranks_reproduce_reference_golden(tmp_path, world):
    out = str(tmp_path / "res")
    port = 31500 + (os.getpid() % 2000) + world
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    g = np.load(GOLDEN)
    res = [torch.load(f"{out}.{r}", weights_only=False) for r in range(world)]
    blocks = [r["rows"] for r in res]
    assert blocks[0][0] == 0 and blocks[-1][1] == 24 and all(a[1] == b[0] for a, b in zip(blocks, blocks[1:]))
    assert all(r["state_cells"] < 24 * 24 for r in res)                 # nobody holds the whole basis
    for r in res:
        assert r["its"] == g["gn_its"].tolist()                          # same decisions on every rank
        assert np.array_equal(r["red"], res[0]["red"])                   # replicated reduced state, bit for bit
        assert np.array_equal(r["snaps"], res[0]["snaps"])
        assert max(rel(r["snaps"][:, i], g["snaps"][:, i]) for i in range(r["snaps"].shape[1])) < TOL
