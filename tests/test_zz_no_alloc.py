"""bfd_gram_static: the normal equations without a workspace argument and without any allocation.

Collected last on purpose (``pytest -x`` stops at the first failure).  The entry point has no
counterpart in the reference: its output is checked against the existing
entry point bfd_gram, through the C ABI, on two calls in a row that share the library's fixed
scratch array; the second call is smaller, so whatever the first call left behind is still there.
"""
import ctypes as C

import numpy as np
import pytest

from finitedifference_b200 import _lib

TOL = 1e-12          # relative to max|G|: the tolerance test_gpu_rom.py uses for bfd_gram against cuBLAS
MAX_ROWS = 1 << 21   # BFD_GRAM_STATIC_MAX_ROWS


@pytest.mark.gpu
def test_gram_static_equals_gram_on_two_calls_in_a_row():
    import torch
    assert torch.cuda.is_available()
    lib = _lib.get_lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    stream = torch.cuda.current_stream().cuda_stream
    gen = torch.Generator(device=dev).manual_seed(20261105)
    # the fine grid's 2N rows with k = 95 first, then a small ragged problem (M not a multiple of 16)
    for M, k in ((2 * 750 * 750, 95), (4001, 37)):
        X = torch.randn((M, k), dtype=torch.float64, device=dev, generator=gen)
        f = torch.randn(M, dtype=torch.float64, device=dev, generator=gen)
        nbytes = lib.bfd_gram_workspace_bytes(M, k)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        G0 = torch.full((k, k), float("nan"), dtype=torch.float64, device=dev)
        b0 = torch.full((k,), float("nan"), dtype=torch.float64, device=dev)
        G1, b1 = G0.clone(), b0.clone()
        _lib.check(lib.bfd_gram(X.data_ptr(), f.data_ptr(), M, k, G0.data_ptr(), b0.data_ptr(), ws.data_ptr(),
                                nbytes, stream), lib, "bfd_gram")
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(dev)
        _lib.check(lib.bfd_gram_static(X.data_ptr(), f.data_ptr(), M, k, G1.data_ptr(), b1.data_ptr(), stream),
                   lib, "bfd_gram_static")
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(dev) == before
        scale = float(G0.abs().max())
        assert float((G1 - G0).abs().max()) < TOL * scale
        assert float((b1 - b0).abs().max()) < TOL * max(float(b0.abs().max()), scale ** 0.5)
        assert torch.equal(G1, G1.T)
        Gr = X.T @ X
        assert float((G1 - Gr).abs().max()) < TOL * scale
    # sizes beyond the fixed scratch's maximum are refused with a reason, nothing is launched
    launches = lib.bfd_launch_count()
    rc = lib.bfd_gram_static(X.data_ptr(), f.data_ptr(), MAX_ROWS + 1, 37, G1.data_ptr(), b1.data_ptr(), stream)
    assert rc == -6 and b"maximum" in lib.bfd_last_cuda_error()
    assert lib.bfd_gram_static(X.data_ptr(), f.data_ptr(), 4001, 97, G1.data_ptr(), b1.data_ptr(), stream) == -6
    assert lib.bfd_launch_count() == launches


def test_gram_static_on_the_emulator():
    """The same two calls on the CPU build of the kernel source (tests/simt_emu)."""
    from tests import emu_driver
    lib = emu_driver.load()
    rng = np.random.default_rng(5)
    for M, k in ((150, 95), (37, 12)):
        X = rng.standard_normal((M, k))
        f = rng.standard_normal(M)
        G, b = np.full((k, k), np.nan), np.full(k, np.nan)
        p = emu_driver.ptr
        _lib.check(lib.bfd_gram_static(p(X), p(f), M, k, p(G), p(b), None), lib, "bfd_gram_static")
        assert np.abs(G - X.T @ X).max() < 1e-12 * np.abs(G).max() and np.array_equal(G, G.T)
        assert np.abs(b - X.T @ f).max() < 1e-12 * np.abs(G).max()
    assert lib.bfd_gram_static(emu_driver.ptr(X), emu_driver.ptr(f), MAX_ROWS + 1, 12, emu_driver.ptr(G),
                               emu_driver.ptr(b), None) == -6
    assert b"maximum" in lib.bfd_last_cuda_error()
