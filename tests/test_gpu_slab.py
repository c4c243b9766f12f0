"""Slab decomposition on the GPU.  The driver's GPU tier has one device, so the slabs of these
tests live in one process on that device (one stream and one host thread per slab); they run the
very kernel, system-scope control path and halo hand-off a multi-GPU run uses -- only the
addresses are local instead of cudaIpc mappings.  tools/bench_slab.py is the multi-process run."""
import numpy as np
import pytest

from oracle import burgers_oracle as orc

pytestmark = pytest.mark.gpu


def run_group(n, steps, mu, world, **kw):
    from finitedifference_b200 import slab as S
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    grp = S.SlabGroup(gx, gy, 0.05, steps, [mu], world, **kw)
    try:
        grp.set_initial(w0)
        grp.run()
        snaps = grp.snapshots(0)
        its, rel = grp.diagnostics()
        redone = grp.slabs[0].steps_redone
    finally:
        grp.close()
    return gx, gy, w0, snaps, its[0], rel[0], redone


@pytest.mark.parametrize("world,n,kw", [(2, 64, {}), (3, 100, {}), (2, 70, dict(newton_hint=1)),
                                        (4, 128, dict(newton_hint=6)), (2, 64, dict(chunk_steps=1))])
def test_slabs_match_oracle(world, n, kw):
    steps, mu = 6, [4.75, 0.02]
    gx, gy, w0, snaps, its, rel, redone = run_group(n, steps, mu, world, **kw)
    ref, hist = orc.fom_trajectory(gx, gy, w0, 0.05, steps, mu, return_info=True)
    assert (np.linalg.norm(snaps - ref, axis=0) / np.linalg.norm(ref, axis=0)).max() < 1e-10
    assert its.tolist() == [len(h) for h in hist]
    if kw.get("newton_hint"):
        assert redone > 0


def test_slabs_equal_single_workspace_run_at_fine_grid():
    """750 x 750, 40 steps: four slabs reproduce the single-GPU solver bit for bit (same kernel,
    same arithmetic order inside a strip; only the norm summation order is shared too)."""
    from finitedifference_b200 import hypernet2D as h
    n, steps, mu = 750, 40, [4.75, 0.02]
    gx, gy, w0, snaps, its, rel, _ = run_group(n, steps, mu, 4)
    one, info = h.inviscid_burgers_implicit2D(gx, gy, w0, 0.05, steps, mu, return_info=True)
    assert np.array_equal(snaps, one)
    assert its.tolist() == list(info["newton_its"])
    assert rel.max() < 1e-12
