"""Host logic of FomSolver.snapshots' staged device-to-host copy (chunk bounds, ragged last chunk,
double buffering) on CPU tensors, with the CUDA pieces (events, pinning) replaced by no-ops; the
device path itself is exercised by every GPU test that asks for host snapshots."""
import contextlib

import numpy as np
import pytest
import torch

from finitedifference_b200.fom import FomSolver, fresh_host_array


class _Event:
    def record(self):
        pass

    def synchronize(self):
        pass


class _Solver:
    device = torch.device("cpu")
    _staged_chunks = FomSolver._staged_chunks


@pytest.mark.parametrize("rows,cols,chunk", [(2, 4, 64 << 20), (1000, 7, 8 * 7 * 33), (1125, 501, 1 << 16), (5, 3, 1)])
def test_staged_copy_chunks(monkeypatch, rows, cols, chunk):
    monkeypatch.setattr(torch.cuda, "Event", _Event)
    monkeypatch.setattr(torch.cuda, "device", lambda d: contextlib.nullcontext())
    monkeypatch.setattr(torch.Tensor, "pin_memory", lambda self: self)
    monkeypatch.setattr(FomSolver, "_PINS", {})
    dev = torch.randn(rows, cols, dtype=torch.float64)
    host = fresh_host_array((rows, cols))
    FomSolver._staged_to_host(_Solver(), dev, host, chunk_bytes=chunk)
    assert np.array_equal(host, dev.numpy())


def test_fresh_host_array_is_plain_numpy():
    a = fresh_host_array((3000, 1500))                      # 36 MB: spans several huge pages
    assert a.dtype == np.float64 and a.flags.c_contiguous and a.flags.writeable and a.shape == (3000, 1500)
    a[:] = 2.0
    assert a.sum() == 2.0 * a.size


def test_save_sink_default_is_the_reference_cache_file(tmp_path):
    """The sink behind compute_snaps_to_cache writes what np.save writes for the reference (hypernet2D.py:3216):
    same name, same (2N, T+1) float64 C-order bytes; the two storage options (float32, every s-th time level) go
    to tagged files the reference's loader does not pick up."""
    import numpy as np
    from finitedifference_b200.hypernet2D import param_to_snap_fn
    from finitedifference_b200.sweep import save_sink
    rng = np.random.default_rng(0)
    snaps = rng.standard_normal((18, 11))
    mu = [4.75, 0.02]
    folder = str(tmp_path / "param_snaps")
    save_sink(folder)(0, mu, snaps)
    fn = param_to_snap_fn(mu, snap_folder=folder)
    ref = tmp_path / "ref.npy"
    np.save(ref, snaps)
    assert open(fn, "rb").read() == open(ref, "rb").read()
    save_sink(folder, dtype=np.float32, time_stride=4)(0, mu, snaps)
    tagged = fn[:-4] + ".every4.f32.npy"
    got = np.load(tagged)
    assert got.dtype == np.float32 and got.flags.c_contiguous
    assert np.array_equal(got, snaps[:, [0, 4, 8, 10]].astype(np.float32))      # the last level is kept
    save_sink(folder, time_stride=5)(0, mu, snaps)
    assert np.array_equal(np.load(fn[:-4] + ".every5.npy"), snaps[:, [0, 5, 10]])
    import pytest
    with pytest.raises(ValueError):
        save_sink(folder, time_stride=0)
    with pytest.raises(ValueError):
        save_sink(folder, dtype=np.int32)
