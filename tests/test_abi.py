"""The C-ABI library builds for sm_100a, loads, and exports every symbol the header declares."""
import ctypes
import os
import re

from finitedifference_b200 import _build, _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "burgers_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bfd_[a-z0-9_]+)\s*\(", text)))


def test_build_and_symbols():
    path = _build.build()
    assert os.path.exists(path)
    lib = ctypes.CDLL(path)          # needs libcudart only, no GPU
    declared = header_symbols()
    assert declared, "no symbols parsed from the header"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/burgers_b200.h but not exported"
    assert sorted(_lib.SIGNATURES) == declared, "ctypes signatures out of sync with the header"
    assert lib.bfd_abi_version() == 5


def test_bad_arguments_are_reported_not_thrown():
    lib = _lib.bind(_build.build())
    d = _lib.fom_desc(0, 10, 1, 1)
    assert lib.bfd_fom_workspace_bytes(ctypes.byref(d)) == 0
    assert lib.bfd_fom_init(ctypes.byref(d), None, 0, None) == -1
    for bad in (dict(newton_hint=_lib.FOM_KMAX + 1), dict(chunk_steps=_lib.FOM_MAX_CHUNK + 1), dict(rel_tol=1.0)):
        assert lib.bfd_fom_workspace_bytes(ctypes.byref(_lib.fom_desc(10, 10, 1, 1, **bad))) == 0
    d = _lib.fom_desc(750, 750, 1, 500)
    nbytes = lib.bfd_fom_workspace_bytes(ctypes.byref(d))
    assert 501 * 2 * 750 * 750 * 8 < nbytes < 1.15 * 501 * 2 * 750 * 750 * 8


def test_no_cpu_fallback_without_gpu():
    import pytest
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import numpy as np
    from finitedifference_b200 import hypernet2D as h
    gx, gy = h.make_2D_grid(0, 100, 0, 100, 8, 8)
    with pytest.raises(_lib.BurgersError):
        h.inviscid_burgers_implicit2D(gx, gy, np.ones(128), 0.05, 1, [4.75, 0.02])


def test_training_matrix_rejects_foreign_callables():
    """res / jac other than the functions the device path evaluates raise instead of being ignored."""
    import pytest
    from finitedifference_b200 import hypernet2D as h
    with pytest.raises(ValueError, match="res="):
        h.compute_ECSW_training_matrix_2D(None, None, None, lambda *a: 0, None)
    with pytest.raises(ValueError, match="jac="):
        h.compute_ECSW_training_matrix_2D(None, None, None, h.inviscid_burgers_res2D, h.inviscid_burgers_res2D)
