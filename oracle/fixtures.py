"""Deterministic inputs shared by oracle/gen_golden_large.py and the tests (test infrastructure).

Nothing here comes from the reference: these are the *inputs* of the large fixtures, built the
same way where the fixture is generated (build container, unmodified reference) and where it is
checked (CPU suite with the oracle, GPU suite with the product).

* ``hat_basis``  an analytic, orthonormalised basis with ones(2N) in its span: reproducible to
  rounding on any machine, so an LSPG fixture at 750^2 with k = 95 needs no 855 MB basis file.
* ``pod_from_coef``  the POD basis U[:, :k] = S W diag(1/sigma) rebuilt from the k columns of
  W/sigma stored in the fixture and snapshots S recomputed by the caller (the FOM is pinned to
  the reference separately, so S is known to ~1e-13).
"""
from __future__ import annotations

import numpy as np


def grid(n, lo=0.0, hi=100.0):
    """make_2D_grid(lo, hi, lo, hi, n, n) of hypernet2D.py:2501-2507."""
    g = np.linspace(lo, hi, n + 1)
    return g, g.copy()


def hat_basis(n, k=95, nodes=24, length=100.0):
    """(2 n^2, k) orthonormal columns: {u, v} x {1, cos(pi y/L)} x `nodes` hat functions in x.

    The hats are a partition of unity, so the reference's projected initial condition
    basis @ (basis.T @ ones) is ones up to rounding (hypernet2D.py:158-160)."""
    assert k <= 4 * nodes
    xc = (np.arange(n) + 0.5) * (length / n)
    yc = xc
    xn = np.linspace(0.0, length, nodes)
    h = xn[1] - xn[0]
    hats = np.maximum(0.0, 1.0 - np.abs(xc[None, :] - xn[:, None]) / h)      # (nodes, n)
    ylv = np.stack([np.ones(n), np.cos(np.pi * yc / length)])                 # (2, n)
    N = n * n
    B = np.zeros((2 * N, 4 * nodes))
    j = 0
    for comp in range(2):
        for lv in range(2):
            for i in range(nodes):
                B[comp * N:(comp + 1) * N, j] = np.outer(ylv[lv], hats[i]).ravel()
                j += 1
    Q, R = np.linalg.qr(B[:, :k])
    # the QR factorisation with a positive diagonal of R is unique: no dependence on LAPACK's
    # Householder sign conventions or on the number of threads
    Q *= np.sign(np.diag(R))[None, :]
    return np.ascontiguousarray(Q)


POD_TIME_STRIDE = 4     # training snapshots: every 4th time level of each training trajectory


def pod_training_matrix(trajs):
    """hstack of every POD_TIME_STRIDE-th snapshot of each (2N, T+1) trajectory."""
    return np.hstack([S[:, ::POD_TIME_STRIDE] for S in trajs])


def pod_from_coef(S, coef):
    """U[:, :k] = S @ coef with coef = W[:, :k] / sigma[:k] (S = U diag(sigma) W^T)."""
    return np.ascontiguousarray(S @ coef)
