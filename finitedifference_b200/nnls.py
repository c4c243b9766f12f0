"""Non-negative least squares for the ECSW weights where the training matrix already is: in HBM
(SURVEY.md 8f item 2).

The reference trains the ECSW weights with ``scipy.optimize.nnls(c, c.sum(axis=1))`` on 20 column
blocks of the training matrix ``C`` (run_HPROM_ecsw_joshua.py:83-86, joblib workers on the host;
``C`` is 4 750 x 562 500 doubles = 21 GB on the fine grid).  This module restates the published
algorithm -- Lawson & Hanson, *Solving Least Squares Problems* (1974), ch. 23, the active-set method
behind both SciPy's ``nnls`` and MATLAB's ``lsqnonneg`` -- on torch tensors, so that the large
products (``A^T A`` once per block, ``A^T (b - A x)`` per iteration) are library GEMM / GEMV calls on
the device; nothing here is a hand-written kernel.  The passive-set least-squares problems are solved
from an incrementally grown Cholesky factor of ``(A^T A)[P, P]`` with a refinement step on the true
residual (minimum-norm solution through ``eigh`` when the columns are dependent).  SciPy is not pinned by the reference (no requirements file); the tests compare with the
SciPy installed in this image (1.18) on problems with a unique solution, and check the KKT conditions
where the solution is not unique (more columns than rows, the reference's case: the active set is
then rounding-sensitive).
"""
from __future__ import annotations

import numpy as np
import torch

from .fom import _require_cuda


class _PassiveFactor:
    """Cholesky factor of the passive columns' Gram matrix, grown one column at a time (O(p^2) per
    entering column); a leaving column triggers a refactorisation.  ``ok`` turns False when the Gram
    matrix is numerically singular (dependent columns): the caller then solves that passive set by
    an eigen-decomposition instead (minimum-norm least squares, what LAPACK's gelsy/gelsd would give)."""

    def __init__(self, AtA):
        self.AtA = AtA
        self.idx = []                   # passive columns in factor order
        self.L = torch.zeros((0, 0), dtype=torch.float64, device=AtA.device)
        self.ok = True

    def _grow(self, j):
        p = len(self.idx)
        g = self.AtA[j, self.idx] if p else self.AtA.new_zeros(0)
        l = torch.linalg.solve_triangular(self.L, g[:, None], upper=False)[:, 0] if p else g
        d2 = float(self.AtA[j, j] - l @ l)
        if not d2 > 1e-14 * float(self.AtA[j, j]):
            self.ok = False
            d2 = max(d2, 0.0)
        Ln = self.AtA.new_zeros((p + 1, p + 1))
        Ln[:p, :p] = self.L
        Ln[p, :p] = l
        Ln[p, p] = d2 ** 0.5
        self.L = Ln
        self.idx.append(j)

    def add(self, j):
        self._grow(int(j))

    def keep(self, mask):
        """Drop the passive columns whose mask entry is False: one Cholesky of what is left."""
        m = mask.cpu().numpy()                                # one copy, not one synchronisation per column
        self.idx = [j for j in self.idx if m[j]]
        self.ok = True
        if not self.idx:
            self.L = self.AtA.new_zeros((0, 0))
            return
        idx = torch.tensor(self.idx, dtype=torch.long, device=self.AtA.device)
        G = self.AtA[idx][:, idx]
        L, info = torch.linalg.cholesky_ex(G)
        if int(info) != 0 or not bool((torch.diagonal(L) ** 2 > 1e-14 * torch.diagonal(G)).all()):
            self.ok = False
        self.L = L


def nnls_tensor(A: torch.Tensor, b: torch.Tensor, maxiter: int | None = None):
    """``argmin ||A x - b||_2, x >= 0`` on A's device; returns ``(x, rnorm, iterations)``.

    Lawson-Hanson.  The passive-set problems are solved from the Cholesky factor of the passive Gram
    matrix followed by a refinement step on the TRUE residual ``A_P^T (b - A_P z)`` (corrected
    semi-normal equations: the error of the plain normal equations, ~eps cond(A_P)^2, is what decides
    the active set when the residual is close to zero -- the reference's case, b = C.sum(axis=1));
    the dual ``w = A^T (b - A x)`` is formed from the residual as well.  Dependent passive columns
    (Cholesky breakdown) are solved by an eigen-decomposition of the Gram block, i.e. the minimum-norm
    solution; a column that would re-enter with a non-positive value is skipped for that iteration
    (Lawson & Hanson's guard against cycling)."""
    if A.ndim != 2 or b.ndim != 1 or A.shape[0] != b.shape[0]:
        raise ValueError(f"incompatible shapes {tuple(A.shape)} and {tuple(b.shape)}")
    A, b = A.to(torch.float64), b.to(torch.float64)
    m, n = A.shape
    dev = A.device
    x = torch.zeros(n, dtype=torch.float64, device=dev)
    if n == 0:
        return x, float(torch.linalg.norm(b)), 0
    maxiter = 3 * n if not maxiter else int(maxiter)
    AtA, Atb = A.T @ A, A.T @ b
    tol = 10.0 * np.finfo(np.float64).eps * float(A.abs().sum(dim=0).max()) * max(m, n)     # lsqnonneg's tolerance
    passive = torch.zeros(n, dtype=torch.bool, device=dev)
    fac = _PassiveFactor(AtA)
    w = Atb.clone()                                      # A^T (b - A x) at x = 0
    its = 0

    def solve_passive():
        idx = torch.tensor(fac.idx, dtype=torch.long, device=dev)
        Ap = A[:, idx]
        if fac.ok:
            z = torch.cholesky_solve(Atb[idx][:, None], fac.L)[:, 0]
            z = z + torch.cholesky_solve((Ap.T @ (b - Ap @ z))[:, None], fac.L)[:, 0]
        else:                                            # dependent columns: minimum-norm least squares
            lam, Q = torch.linalg.eigh(AtA[idx][:, idx])
            keep = lam > lam.max() * len(fac.idx) * np.finfo(np.float64).eps
            z = Q[:, keep] @ ((Q[:, keep].T @ Atb[idx]) / lam[keep])
        s = torch.zeros_like(x)
        s[idx] = z
        return s

    skip = torch.zeros(n, dtype=torch.bool, device=dev)  # columns refused in this outer iteration
    while True:
        cand = torch.where(passive | skip, torch.full_like(w, -np.inf), w)
        j = int(torch.argmax(cand))
        if bool(passive.all()) or float(cand[j]) <= tol:
            break
        if its >= maxiter:
            raise RuntimeError("Maximum number of iterations reached.")
        its += 1
        passive[j] = True
        fac.add(j)
        s = solve_passive()
        if float(s[j]) <= 0.0:                           # the entering column must come in positive
            passive[j] = False
            fac.keep(passive)
            skip[j] = True
            continue
        while bool((s[passive] <= 0).any()):            # step back to the boundary of the feasible set
            bad = torch.nonzero(passive & (s <= 0))[:, 0]
            ratios = x[bad] / (x[bad] - s[bad])
            kmin = int(torch.argmin(ratios))
            x = torch.clamp(x + float(ratios[kmin]) * (s - x), min=0.0)
            x[bad[kmin]] = 0.0                          # the variable that hit its bound leaves the passive set
            passive &= x > 0
            x = torch.where(passive, x, torch.zeros_like(x))
            fac.keep(passive)
            if not bool(passive.any()):
                s = torch.zeros_like(x)
                break
            s = solve_passive()
        x = s
        w = A.T @ (b - A @ x)
        skip.zero_()
    return x, float(torch.linalg.norm(A @ x - b)), its


def nnls(A, b, *, maxiter=None, device=None):
    """``scipy.optimize.nnls`` signature: returns ``(x, rnorm)`` as NumPy / float; runs on the GPU."""
    dev = _require_cuda(device if device is not None else (A.device if isinstance(A, torch.Tensor) and A.is_cuda else None))
    At = A if isinstance(A, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(A, dtype=np.float64))
    bt = b if isinstance(b, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(b, dtype=np.float64).ravel())
    with torch.cuda.device(dev):
        x, rnorm, _ = nnls_tensor(At.to(dev), bt.to(dev), maxiter)
    return x.cpu().numpy(), rnorm


def ecsw_weights(C, blocks=20, *, maxiter=None, device=None):
    """The reference's block-wise training (run_HPROM_ecsw_joshua.py:83-86): ``np.array_split(C, blocks,
    axis=1)``, ``nnls(c, c.sum(axis=1))`` per block, weights concatenated.  ``C``: NumPy array or CUDA
    tensor ``(k*n_snaps, n_cells)``; returns the NumPy weight vector ``(n_cells,)``."""
    dev = _require_cuda(device if device is not None else (C.device if isinstance(C, torch.Tensor) and C.is_cuda else None))
    n = C.shape[1]
    bounds = np.cumsum([0] + [len(p) for p in np.array_split(np.arange(n), blocks)])
    out = []
    with torch.cuda.device(dev):
        for lo, hi in zip(bounds[:-1], bounds[1:]):
            c = C[:, lo:hi]
            c = (c if isinstance(c, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(c, dtype=np.float64))).to(dev)
            x, _, _ = nnls_tensor(c, c.sum(dim=1), maxiter if maxiter else 9999999999)   # the reference's maxiter (:83)
            out.append(x.cpu().numpy())
    return np.hstack(out)
