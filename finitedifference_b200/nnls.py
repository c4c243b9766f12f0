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

Two more of the reference's trainers are compositions over the same pieces: its own ``lsqnonneg``
(BurgersFD_CleanCoarse/lsqnonneg.py:4, MATLAB's active-set method with an absolute tolerance, a relative-error
early stop and a cap on the support) and the three-level block training of
run_HPROM_ecsw_multilevel_complete.py:95-177.
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


def lsqnonneg_tensor(C: torch.Tensor, d: torch.Tensor, x0=None, tol=None, itmax_factor=100, max_support=None,
                     rel_err_thresh=0.01):
    """The reference's ``lsqnonneg`` (BurgersFD_CleanCoarse/lsqnonneg.py:4-109) on C's device: returns
    ``(x, resnorm, residual)`` with ``resnorm = sum(residual**2)``.

    Same rules as the reference's: tolerance ``10 * 2.22e-16 * sum|C| * (max(C.shape) + 1)`` (:20-21; its
    ``norm1`` is the sum of all absolute entries), the entering variable is the largest dual among the zero set
    (:45-47), a passive solution with an entry ``<= tol`` steps back along ``x + alpha (z - x)`` and drops
    ``|x| < tol`` (:66-79), and after every outer iteration the loop stops once ``||resid|| / ||d|| <
    rel_err_thresh`` (default 0.01: the reference's default is an APPROXIMATE fit) or the support has reached
    ``max_support`` (:93-98).  The reference solves each passive set with ``pinv`` of the column-masked matrix
    (:58, :82); here it is the Cholesky factor of the passive Gram block with one refinement step, or its
    eigen-decomposition when the columns are dependent -- the same minimum-norm solution.  ``it > itmax_factor
    * n`` inner iterations raise, as in the reference (:68-71)."""
    if C.ndim != 2 or d.ndim != 1 or C.shape[0] != d.shape[0]:
        raise ValueError(f"incompatible shapes {tuple(C.shape)} and {tuple(d.shape)}")
    C, d = C.to(torch.float64), d.to(torch.float64)
    m, n = C.shape
    dev = C.device
    if tol is None:
        tol = 10.0 * 2.22e-16 * float(C.abs().sum()) * (max(m, n) + 1)
    x = torch.zeros(n, dtype=torch.float64, device=dev)
    if x0 is not None:
        x0 = torch.as_tensor(x0, dtype=torch.float64, device=dev)
        if not bool((x0 < 0).any()):
            x = x0.clone()
    CtC, Ctd = C.T @ C, C.T @ d
    passive = torch.zeros(n, dtype=torch.bool, device=dev)
    fac = _PassiveFactor(CtC)

    def solve_passive():
        z = torch.zeros(n, dtype=torch.float64, device=dev)
        if not fac.idx:
            return z
        idx = torch.tensor(fac.idx, dtype=torch.long, device=dev)
        Cp = C[:, idx]
        if fac.ok:
            zp = torch.cholesky_solve(Ctd[idx][:, None], fac.L)[:, 0]
            zp = zp + torch.cholesky_solve((Cp.T @ (d - Cp @ zp))[:, None], fac.L)[:, 0]
        else:
            lam, Q = torch.linalg.eigh(CtC[idx][:, idx])
            keep = lam > lam.max() * len(fac.idx) * np.finfo(np.float64).eps
            zp = Q[:, keep] @ ((Q[:, keep].T @ Ctd[idx]) / lam[keep])
        z[idx] = zp
        return z

    resid = d - C @ x
    w = C.T @ resid
    it, itmax = 0, int(itmax_factor) * n
    dnorm = float(torch.linalg.norm(d))
    while not bool(passive.all()):
        cand = torch.where(passive, torch.full_like(w, -np.inf), w)
        t = int(torch.argmax(cand))
        if not float(cand[t]) > tol:
            break
        passive[t] = True
        fac.add(t)
        z = solve_passive()
        while bool((z[passive] <= tol).any()):
            it += 1
            if it > itmax:
                raise RuntimeError(f"Iteration limit ({it}) exceeded. Increase tolerance tol "
                                   f"(max_error={float(z[passive].max())}).")
            qq = passive & (z <= tol)
            alpha = float((x[qq] / (x[qq] - z[qq])).min())
            x = x + alpha * (z - x)
            passive = passive & ~(x.abs() < tol)
            fac.keep(passive)
            z = solve_passive()
        x = z
        resid = d - C @ x
        w = C.T @ resid
        rel_err = float(torch.linalg.norm(resid)) / dnorm if dnorm > 0 else 0.0
        if rel_err < rel_err_thresh:
            break
        if max_support is not None and int((x > 0).sum()) >= max_support:
            break
    return x, float((resid * resid).sum()), resid


def lsqnonneg(C, d, x0=None, tol=None, itmax_factor=100, max_support=None, rel_err_thresh=0.01, *, device=None):
    """Drop-in for the reference's ``lsqnonneg(C, d, ...)`` (same arguments, same return triple as NumPy
    arrays / float); the work runs on the GPU."""
    dev = _require_cuda(device if device is not None else (C.device if isinstance(C, torch.Tensor) and C.is_cuda else None))
    Ct = C if isinstance(C, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(C, dtype=np.float64))
    dt = d if isinstance(d, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(d, dtype=np.float64).ravel())
    with torch.cuda.device(dev):
        x, resnorm, resid = lsqnonneg_tensor(Ct.to(dev), dt.to(dev), x0, tol, itmax_factor, max_support, rel_err_thresh)
    return x.cpu().numpy(), resnorm, resid.cpu().numpy()


def multilevel_weights_tensor(C: torch.Tensor, levels=(12, 2), maxiter=None):
    """The three-level training of run_HPROM_ecsw_multilevel_complete.py:95-177 on C's device.

    Level 1 (:95-117): the columns of ``C`` in ``levels[0]`` blocks (``np.array_split``), each block trained
    against its own column sum; the non-zero weights survive.  Level 2 (:119-155): the surviving columns in
    ``levels[1]`` blocks, again each against its own column sum (the script forms ``C_level2 @ z_level2`` at :131
    and does not use it, so from here on the weights integrate the plain sum of the surviving columns, not the
    level-1 quadrature: kept as the script has it).  Last level (:157-177): one NNLS on what is left against ``C_last @ z`` with ``z`` the
    weights of the level before.  Returns ``(weights (n,), rnorm of the last solve, columns alive after each
    level)``.  Any number of block levels is accepted; ``levels=()`` is a single NNLS against the column sum."""
    n = C.shape[1]
    dev = C.device
    maxiter = maxiter if maxiter else 9999999999            # the scripts' maxiter (run_HPROM_ecsw_joshua.py:83)
    alive = torch.arange(n, device=dev)
    z = None
    sizes = []
    for li, nb in enumerate(levels):
        Ck = C if li == 0 else C[:, alive]                   # level 1 works on views of C: no 21 GB gather
        width = Ck.shape[1]
        bounds = np.cumsum([0] + [len(p) for p in np.array_split(np.arange(width), nb)])
        keep, wts = [], []
        for lo, hi in zip(bounds[:-1], bounds[1:]):
            c = Ck[:, int(lo):int(hi)]
            x, _, _ = nnls_tensor(c, c.sum(dim=1), maxiter)
            nz = torch.nonzero(x)[:, 0]
            keep.append(nz + int(lo))
            wts.append(x[nz])
        sel = torch.cat(keep) if keep else torch.zeros(0, dtype=torch.long, device=dev)
        z = torch.cat(wts) if wts else torch.zeros(0, dtype=torch.float64, device=dev)
        alive = alive[sel]
        sizes.append(int(alive.numel()))
    Cl = C[:, alive]
    b = Cl @ z if z is not None else Cl.sum(dim=1)
    w, rnorm, _ = nnls_tensor(Cl, b, maxiter)
    out = torch.zeros(n, dtype=torch.float64, device=dev)
    out[alive] = w
    return out, rnorm, sizes


def ecsw_weights_multilevel(C, levels=(12, 2), *, maxiter=None, device=None):
    """``multilevel_weights_tensor`` for a NumPy array or CUDA tensor ``C (k*n_snaps, n_cells)``: returns the
    NumPy weight vector ``(n_cells,)`` (the script's ``weights_interior``, :171-172)."""
    dev = _require_cuda(device if device is not None else (C.device if isinstance(C, torch.Tensor) and C.is_cuda else None))
    Ct = C if isinstance(C, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(C, dtype=np.float64))
    with torch.cuda.device(dev):
        w, _, _ = multilevel_weights_tensor(Ct.to(dev), levels, maxiter)
    return w.cpu().numpy()
