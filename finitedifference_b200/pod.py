"""POD of snapshot matrices where the snapshots already live: in HBM.

Drop-in for ``POD(snaps, num_modes=None, method='svd', random_state=None) -> (u, s)``
(reference hypernet2D.py:2746-2771; ``run_prom.py:81`` spends 650-890 s there on a
1 125 000 x 4 509 matrix).  Both methods are dense linear algebra -- tall-skinny QR, GEMMs, a small
SVD -- i.e. plain library work: cuSOLVER / cuBLAS through torch, float64; nothing here is a
hand-written kernel and nothing runs on the CPU except drawing the Gaussian test matrix of the
randomized method from NumPy's generator (so that ``random_state`` means what it means in the
reference).

* ``method='svd'``  : thin SVD through a QR factorisation, ``snaps = Q R``, ``R = Ur S Vt``,
  ``u = Q Ur`` -- as accurate as LAPACK's ``gesdd`` on the full matrix (the Gram-matrix shortcut
  would square the condition number).
* ``method='rsvd'`` : the algorithm of ``sklearn.utils.extmath.randomized_svd`` (the reference's
  call, :2767; Halko et al. 2009 with LU-normalised power iterations), restated step by step:
  ``n_oversamples = 10``, ``n_iter = 7`` if ``num_modes < 0.1 min(shape)`` else 4, LU normaliser,
  final QR, SVD of ``Q^T M``, ``svd_flip`` on u.  sklearn is not pinned by the reference (no
  requirements file); the restatement follows scikit-learn 1.9 as installed in this image.
"""
from __future__ import annotations

import numpy as np
import torch

from .fom import _require_cuda


def _svd_flip_u(u, vt):
    """``sklearn.utils.extmath.svd_flip`` (u-based): the largest-magnitude entry of every column
    of u is made positive."""
    idx = torch.argmax(u.abs(), dim=0)
    signs = torch.sign(u[idx, torch.arange(u.shape[1], device=u.device)])
    signs = torch.where(signs == 0, torch.ones_like(signs), signs)
    return u * signs[None, :], vt * signs[:, None]


def _lu_normalise(a):
    """``scipy.linalg.lu(a, permute_l=True)[0]``: P @ L for a tall matrix, without ever forming the
    m x m permutation matrix (at the fine grid m = 1 125 000: ``torch.linalg.lu`` would ask for 9.4 TB)."""
    m, n = a.shape
    LU, piv = torch.linalg.lu_factor(a)
    L = torch.tril(LU[:, :n], diagonal=-1)
    L.diagonal().fill_(1.0)
    perm = np.arange(m)
    for i, p in enumerate(piv.cpu().numpy() - 1):           # LAPACK's sequential row interchanges
        perm[i], perm[p] = perm[p], perm[i]
    out = torch.empty_like(L)
    out[torch.from_numpy(perm).to(a.device)] = L            # row perm[i] of P L is row i of L
    return out


def pod_tensor(S: torch.Tensor, num_modes=None, method="svd", random_state=None, n_oversamples=10, n_iter="auto"):
    """POD of a 2-D float64 tensor on its own device; returns ``(u, s)`` tensors."""
    if S.ndim != 2:
        raise ValueError("snapshot matrix must be 2-D")
    S = S.to(torch.float64)
    m, n = S.shape
    if method == "svd":
        if m >= n:
            q, r = torch.linalg.qr(S, mode="reduced")
            ur, s, _ = torch.linalg.svd(r, full_matrices=False)
            return q @ ur, s
        u, s, _ = torch.linalg.svd(S, full_matrices=False)
        return u, s
    if method != "rsvd":
        raise ValueError("Unknown method '{}' for POD. Use 'svd' or 'rsvd'.".format(method))
    if num_modes is None:
        num_modes = min(m, n)
    from sklearn.utils import check_random_state           # NumPy RandomState semantics of the reference's call
    rng = check_random_state(random_state)
    n_random = num_modes + n_oversamples
    if n_iter == "auto":
        n_iter = 7 if num_modes < 0.1 * min(m, n) else 4
    M = S.T if m < n else S                                 # transpose='auto'
    Q = torch.from_numpy(rng.normal(size=(M.shape[1], n_random))).to(S.device)
    for _ in range(n_iter):                                 # LU-normalised power iterations (n_iter > 2)
        Q = _lu_normalise(M @ Q) if n_iter > 2 else M @ Q
        Q = _lu_normalise(M.T @ Q) if n_iter > 2 else M.T @ Q
    Q, _ = torch.linalg.qr(M @ Q, mode="reduced")
    B = Q.T @ M
    uhat, s, vt = torch.linalg.svd(B, full_matrices=False)
    u = Q @ uhat
    if m < n:                                               # undo the transposition; flip on v (= our u)
        u, vt = vt.T, u.T
        vt2, u2 = _svd_flip_u(u.clone(), vt.clone())        # decision on the rows of the transposed problem's v
        u, vt = vt2, u2
        return u[:, :num_modes].contiguous(), s[:num_modes]
    u, vt = _svd_flip_u(u, vt)
    return u[:, :num_modes].contiguous(), s[:num_modes]


def POD(snaps, num_modes=None, method="svd", random_state=None, *, device=None, out_device=False):
    """Reference signature (hypernet2D.py:2746).  ``snaps``: NumPy array or CUDA tensor
    ``(2N, n_snaps)``; returns NumPy ``(u, s)`` like the reference, or tensors with ``out_device``."""
    dev = _require_cuda(device if device is not None else (snaps.device if isinstance(snaps, torch.Tensor) and
                                                           snaps.is_cuda else None))
    S = snaps if isinstance(snaps, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(snaps, dtype=np.float64))
    with torch.cuda.device(dev):
        u, s = pod_tensor(S.to(dev), num_modes, method, random_state)
    if out_device:
        return u, s
    return u.cpu().numpy(), s.cpu().numpy()
