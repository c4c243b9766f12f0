"""Slab decomposition of one grid on CPU: the product's kernel compiled against
the SIMT emulator, two or three "ranks" running concurrently on host threads of one process, wired
together exactly as the GPU ranks are (rank 0's control blocks, the next rank's halo buffer) and
checked against the oracle and the single-workspace path."""
import numpy as np
import pytest

from finitedifference_b200 import slab as S
from oracle import burgers_oracle as orc
from tests import emu_driver as E


@pytest.fixture(scope="module")
def lib():
    return E.load()


@pytest.mark.parametrize("nx,world", [(64, 2), (65, 3), (250, 8), (8192, 8), (33, 2)])
def test_partition_is_contiguous_and_balanced(nx, world):
    parts = S.slab_partition(nx, world)
    ns = (nx + 31) // 32
    assert parts[0][0] == 0 and parts[-1][1] == ns
    assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
    sizes = [e - b for b, e in parts]
    assert min(sizes) >= 1 and max(sizes) - min(sizes) <= 1
    cols = S.slab_columns(nx, world)
    assert cols[0][0] == 0 and cols[-1][1] == nx and sum(c1 - c0 for c0, c1 in cols) == nx


def test_partition_needs_a_strip_per_rank():
    with pytest.raises(ValueError):
        S.slab_partition(40, 3)


def run_group(lib, n_x, n_y, steps, mu, world, **kw):
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, n_x, n_y)
    w0 = np.ones(2 * n_x * n_y)
    grp = S.SlabGroup(gx, gy, 0.05, steps, [mu], world, lib=lib, host_memory=True, **kw)
    try:
        grp.set_initial(w0)
        grp.run()
        snaps = grp.snapshots(0)
        its, rel = grp.diagnostics()
        stats = [(s.launches, s.steps_redone) for s in grp.slabs]
    finally:
        grp.close()
    ref, hist = orc.fom_trajectory(gx, gy, w0, 0.05, steps, mu, return_info=True)
    return snaps, its[0], ref, [len(h) for h in hist], stats


@pytest.mark.timeout(600)
@pytest.mark.parametrize("world,n_x,kw", [
    (2, 64, {}),                                        # one strip per rank
    (2, 72, {}),                                        # ragged last strip on rank 1
    (3, 96, dict(newton_hint=1)),                       # under-speculation: continuation passes across ranks
    (2, 64, dict(newton_hint=6)),                       # over-speculation: strict redo across ranks
    (2, 64, dict(chunk_steps=1)),                       # one step per launch: the launch hand-shake every step
    (2, 64, dict(n_y=60)),                              # tall enough for the unmasked main phases, records crossing ranks
    (2, 72, dict(kernel="tile")),                       # fom_tile_kernel across ranks (peer records, system-scope control)
    (3, 96, dict(kernel="tile", newton_hint=1)),
])
def test_slabs_match_oracle(lib, world, n_x, kw):
    kw = dict(kw)
    n_y, steps, mu = kw.pop("n_y", 12), 3, [4.75, 0.02]
    snaps, its, ref, ref_its, stats = run_group(lib, n_x, n_y, steps, mu, world, **kw)
    err = (np.linalg.norm(snaps - ref, axis=0) / np.linalg.norm(ref, axis=0)).max()
    assert err < 1e-10, err
    assert its.tolist() == ref_its
    if kw.get("newton_hint") in (1, 6):
        assert stats[0][1] > 0                          # the speculation really was wrong
