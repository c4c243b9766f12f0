"""SnapshotSweep (groups of samples on worker streams, host output overlapped with the next
group's sweeps): every trajectory equals the one the single-sample drop-in call returns, bit for
bit, in all three output modes; cache files are interchangeable with the reference's."""
import numpy as np
import pytest

from oracle import burgers_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def case():
    from finitedifference_b200 import hypernet2D as h
    n, steps = 40, 7
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    mus = [[m1, m2] for m1 in np.linspace(4.25, 5.5, 3) for m2 in np.linspace(0.015, 0.03, 3)][:7]
    w0 = np.ones(2 * n * n)
    single = [h.inviscid_burgers_implicit2D(gx, gy, w0, 0.05, steps, mu) for mu in mus]
    return h, gx, gy, w0, steps, mus, single


@pytest.mark.parametrize("group,workers", [(2, 3), (3, 2), (1, 1), (4, 4)])
def test_sweep_sink_mode(case, group, workers):
    from finitedifference_b200.sweep import SnapshotSweep
    h, gx, gy, w0, steps, mus, single = case
    got = {}
    sweep = SnapshotSweep(gx, gy, 0.05, steps, group=group, workers=workers, pool=2)
    for _ in range(2):                          # solvers, staging and pool are reused by the second run
        got.clear()
        assert sweep.run(w0, mus, sink=lambda i, mu, s: got.__setitem__(i, (list(mu), s.copy()))) is None
        assert sorted(got) == list(range(len(mus)))
        for i, (mu, s) in got.items():
            assert mu == mus[i] and s.shape == single[i].shape
            assert np.array_equal(s, single[i])
    assert sweep.launches > 0
    assert np.array_equal(w0, np.ones_like(w0))


def test_sweep_out_mode_and_oracle(case):
    from finitedifference_b200.sweep import SnapshotSweep
    h, gx, gy, w0, steps, mus, single = case
    sweep = SnapshotSweep(gx, gy, 0.05, steps, group=2, workers=2)
    out = [sweep.host_buffer() if i % 2 else np.empty((w0.size, steps + 1)) for i in range(len(mus))]
    res = sweep.run(w0, mus, out=out)
    for i, s in enumerate(res):
        assert isinstance(s, np.ndarray) and s.flags.c_contiguous and np.array_equal(s, single[i])
    ref = orc.fom_trajectory(gx, gy, w0, 0.05, steps, mus[5])
    assert (np.linalg.norm(res[5] - ref, axis=0) / np.linalg.norm(ref, axis=0)).max() < 1e-10
    fresh = sweep.run(w0, mus[:2])
    assert all(np.array_equal(a, b) for a, b in zip(fresh, single))
    with pytest.raises(ValueError):
        sweep.run(w0, mus, out=out[:2])
    with pytest.raises(ValueError):
        sweep.run(w0[:-1], mus)


def test_sweep_errors_surface(case):
    from finitedifference_b200.sweep import SnapshotSweep
    h, gx, gy, w0, steps, mus, single = case

    def bad_sink(i, mu, s):
        raise RuntimeError("sink failed")
    with pytest.raises(RuntimeError, match="sink failed"):
        SnapshotSweep(gx, gy, 0.05, steps).run(w0, mus, sink=bad_sink)


def test_cache_files_interchangeable(case, tmp_path):
    h, gx, gy, w0, steps, mus, single = case
    folder = str(tmp_path / "param_snaps")
    names = h.compute_snaps_to_cache(mus[:5], gx, gy, w0, 0.05, steps, snap_folder=folder, group=2, workers=2)
    assert names == [orc.param_to_snap_fn(mu, snap_folder=folder) for mu in mus[:5]]
    for fn, s in zip(names, single):
        assert np.array_equal(np.load(fn), s)
    # a second call finds them; the single-sample loader reads the same files
    assert h.compute_snaps_to_cache(mus[:5], gx, gy, w0, 0.05, steps, snap_folder=folder) == names
    assert np.array_equal(h.load_or_compute_snaps(mus[3], gx, gy, w0, 0.05, steps, snap_folder=folder), single[3])
