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
