"""Drive the C ABI with host (NumPy) buffers against the SIMT-emulated library (CPU tests)."""
import ctypes as C

import numpy as np

from finitedifference_b200 import _lib
from finitedifference_b200.discretisation import is_uniform, parameter_terms, stencil_coefficients
from tests.simt_emu.build_emu import build


def load():
    return _lib.bind(build())


def ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def fom_run(lib, grid_x, grid_y, w0, dt, num_steps, mus, max_its=100, rel_tol=1e-12, chunk_steps=0,
            newton_hint=0, strict=True, uniform_y=None, stats=None, kernel=0):
    """Returns (snaps[batch, 2N, T+1], its[batch, T], relnorm[batch, T])."""
    mus = np.atleast_2d(np.asarray(mus, dtype=np.float64))
    nx, ny, B = grid_x.size - 1, grid_y.size - 1, mus.shape[0]
    coefs = [np.ascontiguousarray(a) for a in stencil_coefficients(grid_x, grid_y, dt)]
    if uniform_y is None:
        uniform_y = is_uniform(coefs[2], coefs[3])
    d = _lib.fom_desc(nx, ny, B, num_steps, max_its, rel_tol, chunk_steps, newton_hint, strict, uniform_y, kernel)
    nbytes = lib.bfd_fom_workspace_bytes(C.byref(d))
    assert nbytes > 0
    ws = np.zeros(nbytes // 8 + 1, dtype=np.float64)
    src, lbc = parameter_terms(grid_x, grid_y, dt, mus)
    coef = _lib.FomCoef(*[ptr(a) for a in coefs], ptr(src), ptr(lbc))
    _lib.check(lib.bfd_fom_init(C.byref(d), ptr(ws), nbytes, None), lib, "init")
    w0 = np.ascontiguousarray(w0, dtype=np.float64)
    stride = 0 if w0.ndim == 1 else w0.shape[1]
    _lib.check(lib.bfd_fom_set_initial(C.byref(d), ptr(ws), ptr(w0), stride, None), lib, "set_initial")
    pinned = np.zeros(16, dtype=np.int32)
    _lib.check(lib.bfd_fom_run(C.byref(d), ptr(ws), C.byref(coef), ptr(pinned), None), lib, "run")
    assert pinned[0] == B and pinned[1] == 0
    if stats is not None:
        stats.update(launches=int(pinned[2]), redone=int(pinned[3]))
    snaps = np.zeros((B, 2 * nx * ny, num_steps + 1))
    snaps_t = np.zeros((B, num_steps + 1, 2 * nx * ny))
    for m in range(B):
        _lib.check(lib.bfd_fom_export(C.byref(d), ptr(ws), m, 0, num_steps + 1, 1, num_steps + 1,
                                      ptr(snaps[m]), None), lib, "export")
        _lib.check(lib.bfd_fom_export(C.byref(d), ptr(ws), m, 0, num_steps + 1, 0, 0,
                                      ptr(snaps_t[m]), None), lib, "export")
        assert np.array_equal(snaps_t[m].T, snaps[m])
    its = np.zeros((B, num_steps), dtype=np.int32)
    rel = np.zeros((B, num_steps))
    _lib.check(lib.bfd_fom_diagnostics(C.byref(d), ptr(ws), ptr(its), ptr(rel), None), lib, "diag")
    return snaps, its, rel
