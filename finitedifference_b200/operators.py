"""The reference's operator-level callables, evaluated by the device kernels.

``run_HPROM_ecsw_joshua.py:61-64`` and the manifold-ROM scripts pass
``inviscid_burgers_res2D`` / ``inviscid_burgers_exact_jac2D`` around as callables and build
operators with ``get_ops``; this module keeps those names and positional signatures
(reference ``BurgersFD_CleanFine/hypernet2D.py``):

    make_ddx :2486, get_ops :2509, inviscid_burgers_res2D :2544, _res2D_alt :2588,
    _res2D_ecsw :2648, inviscid_burgers_exact_jac2D :2703, _exact_jac2D_ecsw :2734.

Residuals run ``bfd_rom_residual`` on index tables derived from the grids; Jacobians run
``bfd_jac_csr_values`` on index tables derived from the OPERATORS THE CALLER PASSES (any
backward-difference pair with one positive and at most one negative entry per row, so both
``get_ops``' operator and the rows-only-permuted one of :167 work) and are returned as SciPy CSR
matrices.  Inputs and outputs are NumPy / SciPy on the host, as in the reference; the trajectory
solvers in ``finitedifference_b200.hypernet2D`` do not go through these wrappers.
"""
from __future__ import annotations

import ctypes as C
from collections import OrderedDict

import numpy as np
import scipy.sparse as sp
import torch

from . import _lib
from .fom import _require_cuda
from .rom import MeshTables

_CACHE: "OrderedDict[tuple, MeshTables]" = OrderedDict()
_CACHE_SLOTS = 8


def _cached(key, build):
    if key in _CACHE:
        _CACHE.move_to_end(key)
        return _CACHE[key]
    mesh = build()
    _CACHE[key] = mesh
    while len(_CACHE) > _CACHE_SLOTS:
        _CACHE.popitem(last=False)
    return mesh


def _digest(*arrays):
    return tuple(hash(np.ascontiguousarray(a).tobytes()) for a in arrays)


def _host(a):
    if torch.is_tensor(a):                                     # :2652-2655 accepts tensors
        a = a.detach().squeeze().cpu().numpy()
    return np.ascontiguousarray(a, dtype=np.float64).ravel()


# ---------------------------------------------------------------------------------------------
# operators (host, one-off)
# ---------------------------------------------------------------------------------------------
def make_ddx(grid_x):
    """Backward difference ``A[i,i] = 1/dx[i]``, ``A[i,i-1] = -1/dx[i-1]`` (:2486-2492), LIL."""
    g = np.asarray(grid_x, dtype=np.float64)
    n = g.size - 1
    inv = 1.0 / (g[1:] - g[:-1])
    return sp.diags([-inv[:-1], inv], [-1, 0], shape=(n, n), format="lil")


def get_ops(grid_x, grid_y):
    """``(Dxec, Dyec, JDxec, JDyec, Eye)`` of :2509-2520: cells numbered ``r*nx + c``; ``JDxec``
    differences along x inside a row, ``JDyec`` along y between rows (built directly, which is
    what the reference's kron followed by the row-and-column permutation amounts to)."""
    gx, gy = np.asarray(grid_x, dtype=np.float64), np.asarray(grid_y, dtype=np.float64)
    nx, ny = gx.size - 1, gy.size - 1
    n = nx * ny
    ix, iy = 1.0 / (gx[1:] - gx[:-1]), 1.0 / (gy[1:] - gy[:-1])
    r, c = np.divmod(np.arange(n), nx)
    jdx = sp.csr_matrix((np.concatenate((ix[c], -ix[c[c >= 1] - 1])),
                         (np.concatenate((np.arange(n), np.arange(n)[c >= 1])),
                          np.concatenate((np.arange(n), np.arange(n)[c >= 1] - 1)))), shape=(n, n))
    jdy = sp.csr_matrix((np.concatenate((iy[r], -iy[r[r >= 1] - 1])),
                         (np.concatenate((np.arange(n), np.arange(n)[r >= 1])),
                          np.concatenate((np.arange(n), np.arange(n)[r >= 1] - nx)))), shape=(n, n))
    return make_ddx(gx), make_ddx(gy), jdx, jdy, sp.identity(2 * n)


# ---------------------------------------------------------------------------------------------
# residuals
# ---------------------------------------------------------------------------------------------
def _residual(mesh, w, wp, device):
    lib = _lib.get_lib()
    wd = torch.from_numpy(w).to(device)
    wpd = torch.from_numpy(wp).to(device)
    r = torch.empty(2 * mesh.nrows, dtype=torch.float64, device=device)
    _lib.check(lib.bfd_rom_residual(C.byref(mesh.c), wd.data_ptr(), wpd.data_ptr(), r.data_ptr(), None,
                                    torch.cuda.current_stream(device).cuda_stream), lib, "bfd_rom_residual")
    return r.cpu().numpy()


def inviscid_burgers_res2D(w, grid_x, grid_y, dt, wp, mu, Dxec=None, Dyec=None, *, device=None):
    """Residual of one trapezoidal step, ``(2N,)`` (:2544-2586).  ``Dxec``/``Dyec`` are implied by
    the grids and ignored."""
    device = _require_cuda(device)
    key = ("full", device.index, float(dt), float(mu[0]), float(mu[1])) + _digest(grid_x, grid_y)
    mesh = _cached(key, lambda: MeshTables(grid_x, grid_y, dt, list(mu)[:2], device))
    return _residual(mesh, _host(w), _host(wp), device)


def inviscid_burgers_res2D_alt(w, grid_x, grid_y, dt, wp, mu, JDxec=None, JDyec=None, *, device=None):
    """Same residual through the lifted operators (:2588-2646); identical values."""
    return inviscid_burgers_res2D(w, grid_x, grid_y, dt, wp, mu, device=device)


def inviscid_burgers_res2D_ecsw(w, grid_x, grid_y, dt, wp, mu, JDxec, JDyec, sample_inds, augmented_sample,
                                lbc=None, src=None, *, device=None):
    """Residual rows at the sampled cells, ``(2 n_s,)`` (:2648-2701).  ``w``/``wp`` are either the
    full state or its restriction ``w[[aug, N + aug]]``, as in the reference; ``lbc`` / ``src`` (one value
    per sampled cell) replace the terms derived from the grids when given."""
    device = _require_cuda(device)
    s = np.asarray(sample_inds, dtype=np.int64)
    a = np.asarray(augmented_sample, dtype=np.int64)
    key = ("ecsw", device.index, float(dt), float(mu[0]), float(mu[1])) + _digest(grid_x, grid_y, s, a)
    mesh = _cached(key, lambda: MeshTables(grid_x, grid_y, dt, list(mu)[:2], device, rows=s, state_cells=a))
    if lbc is not None or src is not None:                                     # precomputed terms (:2658-2669), as the
        mesh = mesh.with_terms(src=src, lbc=lbc)                               # manifold drivers pass them (:652, :856, :1081)
    idx = np.concatenate((a, (np.asarray(grid_x).size - 1) * (np.asarray(grid_y).size - 1) + a))
    w, wp = _host(w), _host(wp)
    w = w[idx] if w.size > 2 * a.size else w                                   # :2672
    wp = wp[idx] if wp.size > 2 * a.size else wp
    return _residual(mesh, np.ascontiguousarray(w), np.ascontiguousarray(wp), device)


# ---------------------------------------------------------------------------------------------
# Jacobians
# ---------------------------------------------------------------------------------------------
def _split_rows(op, what):
    """Per row of a backward-difference operator: (col, value) of its positive entry and of its
    (optional) negative entry."""
    coo = sp.coo_matrix(op)
    coo.sum_duplicates()
    keep = coo.data != 0
    row, col, val = coo.row[keep], coo.col[keep], coo.data[keep]
    n = op.shape[0]
    pos, neg = val > 0, val < 0
    if (np.bincount(row[pos], minlength=n) != 1).any() or (np.bincount(row[neg], minlength=n) > 1).any():
        raise ValueError(f"{what}: expected one positive and at most one negative entry per row")
    pc, pv = np.full(n, -1, dtype=np.int64), np.zeros(n)
    nc, nv = np.full(n, -1, dtype=np.int64), np.zeros(n)
    pc[row[pos]], pv[row[pos]] = col[pos], val[pos]
    nc[row[neg]], nv[row[neg]] = col[neg], val[neg]
    return pc, pv, nc, nv


def jacobian_tables(JDxec, JDyec, dt, own):
    """``bfd_rom_mesh`` tables read off a pair of backward-difference operators: the x-difference's
    positive entry must sit on the row's own cell ``own[i]``; the y-difference's two entries may sit
    anywhere (hypernet2D.py:167 puts them on the transposed cells)."""
    xp, xv, xn, xnv = _split_rows(JDxec, "JDxec")
    yp, yv, yn, ynv = _split_rows(JDyec, "JDyec")
    if not np.array_equal(xp, own):
        raise ValueError("JDxec's positive entry must sit on the row's own cell")
    z = np.zeros(JDxec.shape[0])
    return dict(p=np.asarray(own, dtype=np.int64), xl=xn, yd=yn, ja=yp, jb=yn, cx=0.5 * dt * xv, cxl=-0.5 * dt * xnv,
                cy=0.5 * dt * yv, cyd=-0.5 * dt * ynv, src=z, lbc=z)


def assemble_csr(vals, t, ncells):
    """CSR ``(2 nrows, 2 ncells)`` from the 12 value slots per row of ``bfd_jac_csr_values``."""
    p = t["p"]
    nrows, N = p.size, ncells
    sub = lambda a: np.where(a >= 0, a, p)                     # absent neighbour: its value slot holds 0
    xl, ja, jb = sub(t["xl"]), sub(t["ja"]), sub(t["jb"])
    cols = np.stack((p, xl, ja, jb, N + ja, N + jb, p, xl, N + p, N + xl, N + ja, N + jb), axis=1)
    i = np.arange(nrows)[:, None]
    rows = np.concatenate((np.broadcast_to(i, (nrows, 6)), np.broadcast_to(nrows + i, (nrows, 6))), axis=1)
    J = sp.coo_matrix((np.asarray(vals).ravel(), (rows.ravel(), cols.ravel())),
                      shape=(2 * nrows, 2 * N)).tocsr()          # duplicates (ja == p) are summed
    J.eliminate_zeros()
    return J


def _jacobian(w, dt, JDxec, JDyec, own, device):
    nrows, ncells = JDxec.shape
    if JDyec.shape != (nrows, ncells):
        raise ValueError("JDxec and JDyec differ in shape")
    JDxec, JDyec = sp.csr_matrix(JDxec), sp.csr_matrix(JDyec)
    key = ("ops", device.index, float(dt), nrows, ncells) + \
        _digest(own, JDxec.data, JDxec.indices, JDxec.indptr, JDyec.data, JDyec.indices, JDyec.indptr)

    def build():
        t = jacobian_tables(JDxec, JDyec, dt, own)
        m = MeshTables.from_tables(t, ncells, device)
        m.host = t
        return m

    mesh = _cached(key, build)
    lib = _lib.get_lib()
    wd = torch.from_numpy(w).to(device)
    vals = torch.empty((nrows, 12), dtype=torch.float64, device=device)
    _lib.check(lib.bfd_jac_csr_values(C.byref(mesh.c), wd.data_ptr(), vals.data_ptr(),
                                      torch.cuda.current_stream(device).cuda_stream), lib, "bfd_jac_csr_values")
    return assemble_csr(vals.cpu().numpy(), mesh.host, ncells)


def inviscid_burgers_exact_jac2D(w, dt, JDxec, JDyec, Eye=None, *, device=None):
    """``bmat([[ul, ur], [ll, lr]]) + Eye`` of :2703-2732 as CSR ``(2N, 2N)``; the identity is
    added by the kernel (``Eye`` is accepted for signature compatibility)."""
    device = _require_cuda(device)
    return _jacobian(_host(w), float(dt), JDxec, JDyec, np.arange(JDxec.shape[0], dtype=np.int64), device)


def inviscid_burgers_exact_jac2D_ecsw(w, dt, JDxec, JDyec, Eye, sample_inds, augmented_inds, *, device=None):
    """Sampled rows x augmented columns of the Jacobian, CSR ``(2 n_s, 2 n_aug)`` (:2734-2744).
    ``w`` is the full state or its restriction to the augmented cells."""
    device = _require_cuda(device)
    s = np.asarray(sample_inds, dtype=np.int64)
    a = np.asarray(augmented_inds, dtype=np.int64)
    own = np.searchsorted(a, s)
    if own.size and (own.max() >= a.size or not np.array_equal(a[own], s)):
        raise ValueError("sample_inds is not a subset of augmented_inds")
    w = _host(w)
    if w.size > 2 * a.size:                                                     # :2736
        w = np.concatenate((w[: w.size // 2][a], w[w.size // 2:][a]))
    return _jacobian(np.ascontiguousarray(w), float(dt), JDxec, JDyec, own, device)
