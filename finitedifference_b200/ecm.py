"""Empirical Cubature Method on torch tensors.

Restates ``EmpiricalCubatureMethod`` (reference BurgersFD_CleanFine/empirical_cubature_method.py:15-294;
Hernandez 2020, "A multiscale method for periodic structures using domain decomposition and
ECM-hyperreduction") -- the element selection ``run_HPROM_ecm.py:83-91`` uses instead of NNLS -- with the same
constructor / ``SetUp`` / ``Run`` interface and result attributes ``z`` (selected columns, in selection
order) and ``w`` (their weights).  The matrix ``G`` (basis of the projected residuals, one column per
candidate cell) stays on the device it is given on; every iteration is one GEMV over all candidates
(``G^T r``), an argmax and a rank-one update of ``(G_z^T G_z)^-1`` -- library calls, no hand-written kernel.

Covered: ``InitialCandidatesSet=None`` (what the reference's drivers pass), ``constrain_sum_of_weights``
with and without ``constrain_conditions``, ``Filter_tolerance``.  A candidate set supplied by the caller
(the complement-expansion path, :110-135, :138-142) is not used by any Burgers driver and raises.
"""
from __future__ import annotations

import numpy as np
import torch


class EmpiricalCubatureMethod:
    def __init__(self, ECM_tolerance=0, Filter_tolerance=0, Plotting=False, MaximumNumberUnsuccesfulIterations=100):
        self.ECM_tolerance = ECM_tolerance
        self.Filter_tolerance = Filter_tolerance
        self.Name = "EmpiricalCubature"
        self.Plotting = Plotting
        self.MaximumNumberUnsuccesfulIterations = MaximumNumberUnsuccesfulIterations

    # -- :39-86 ------------------------------------------------------------------------------------
    def SetUp(self, ResidualsBasis, InitialCandidatesSet=None, constrain_sum_of_weights=True,
              constrain_conditions=False, number_of_conditions=0, device=None):
        if InitialCandidatesSet is not None:
            raise ValueError("EmpiricalCubatureMethod: a caller-supplied candidate set is not supported "
                             "(no Burgers driver of the reference passes one)")
        B = ResidualsBasis if isinstance(ResidualsBasis, torch.Tensor) else \
            torch.from_numpy(np.ascontiguousarray(ResidualsBasis, dtype=np.float64))
        if device is not None:
            B = B.to(device)
        B = B.to(torch.float64)
        self.device = B.device
        M = B.shape[0]
        self.W = torch.ones(M, dtype=torch.float64, device=self.device)
        G = B.T.contiguous()
        self.add_constrain_count = None
        elements = torch.ones(M, dtype=torch.float64, device=self.device)
        begin = M - number_of_conditions
        elements[begin:] = 0

        def project(v, G):
            p = v - G.T @ (G @ v)
            return p / torch.linalg.norm(p)

        if constrain_sum_of_weights and not constrain_conditions:
            G = torch.vstack([G, project(elements, G)[None, :]])
            self.add_constrain_count = -1
        elif constrain_sum_of_weights and constrain_conditions:
            G = torch.vstack([G, project(elements, G)[None, :]])
            conditions = torch.ones(M, dtype=torch.float64, device=self.device)
            conditions[:begin] = 0
            G = torch.vstack([G, project(conditions, G)[None, :]])
            self.add_constrain_count = -2
        self.G = G.contiguous()
        self.b = self.G @ self.W
        self.UnsuccesfulIterations = 0

    # -- :88-131 -----------------------------------------------------------------------------------
    def Initialize(self):
        G = self.G
        self.GnormNOONE = torch.linalg.norm(G[:self.add_constrain_count, :] if self.add_constrain_count else G, dim=0)
        M = G.shape[1]
        normB = float(torch.linalg.norm(self.b))
        self.candidate = torch.ones(M, dtype=torch.bool, device=self.device)          # the set y as a mask
        if self.Filter_tolerance > 0:
            self.candidate &= ~(self.GnormNOONE < self.Filter_tolerance * normB)
        self.z = np.zeros(0, dtype=np.int64)
        self.mPOS = 0
        self.r = self.b.clone()
        self.m = self.b.numel()
        self.nerror = float(torch.linalg.norm(self.r)) / normB
        self.nerrorACTUAL = self.nerror

    def Run(self):
        self.Initialize()
        self.Calculate()

    # -- :144-250 ----------------------------------------------------------------------------------
    def Calculate(self):
        G, b = self.G, self.b
        normB = float(torch.linalg.norm(b))
        MaximumLengthZ = 0
        k = 1
        self.success = True
        H = alpha = None
        z = []                                             # selected columns, selection order
        neg_inf = torch.tensor(-np.inf, dtype=torch.float64, device=self.device)
        while self.nerrorACTUAL > self.ECM_tolerance and self.mPOS < self.m and bool(self.candidate.any()):
            # step 1: the candidate most aligned with the residual (:160-165)
            obj = torch.where(self.candidate, G.T @ self.r, neg_inf)
            i = int(torch.argmax(obj))
            gi = G[:, i]
            if k == 1:
                alpha = (gi @ b / (gi @ gi)).reshape(1, 1)           # lstsq(G[:, [i]], b) (:167)
                H = (1.0 / (gi @ gi)).reshape(1, 1)
            else:
                H, alpha = self._update_weights_inverse(G[:, z], H, gi, alpha)
            z.append(i)                                    # step 3 (:172-187)
            n_cand = int(self.candidate.sum())
            if n_cand == 1:
                self.success = False
                break
            self.candidate[i] = False
            # step 4: negative weights go back to the candidates (:189-196)
            if bool((alpha < 0).any()):
                neg = torch.nonzero(alpha[:, 0] <= 0.0)[:, 0].cpu().numpy()
                back = [z[q] for q in neg]
                self.candidate[torch.tensor(back, dtype=torch.long, device=self.device)] = True
                z = [c for q, c in enumerate(z) if q not in set(neg.tolist())]
                H = self._multi_update_inverse_hermitian(H, neg)
                alpha = (H @ (G[:, z].T @ b)).reshape(-1, 1)
            if len(z) > MaximumLengthZ:
                self.UnsuccesfulIterations = 0
            else:
                self.UnsuccesfulIterations += 1
            # step 6: residual (:203-211)
            self.r = b - (G[:, z] @ alpha)[:, 0]
            self.nerror = float(torch.linalg.norm(self.r)) / normB
            self.nerrorACTUAL = self.nerror
            self.mPOS = len(z)
            MaximumLengthZ = max(MaximumLengthZ, len(z))
            k += 1
        self.z = np.asarray(z, dtype=np.int64)
        self.w = (alpha.T * torch.sqrt(self.W[torch.tensor(z, dtype=torch.long, device=self.device)])).cpu().numpy()
        self.iterations = k

    # -- :252-268 ----------------------------------------------------------------------------------
    def _update_weights_inverse(self, A, Aast, a, xold):
        c = A.T @ a
        d = (Aast @ c).reshape(-1, 1)
        s = a @ a - (c.reshape(1, -1) @ d)[0, 0]
        top = torch.hstack([Aast + (d @ d.T) / s, -d / s])
        bottom = torch.hstack([(-d.T / s)[0], (1.0 / s).reshape(1)])
        Bast = torch.vstack([top, bottom[None, :]])
        v = (a @ self.r) / s
        x = torch.vstack([xold - d * v, v.reshape(1, 1)])
        return Bast, x

    # -- :270-294 ----------------------------------------------------------------------------------
    def _multi_update_inverse_hermitian(self, invH, neg_indexes):
        neg_indexes = np.sort(neg_indexes)
        for i in range(neg_indexes.size):
            invH = self._update_inverse_hermitian(invH, int(neg_indexes[i]) - i)
        return invH

    @staticmethod
    def _update_inverse_hermitian(invH, q):
        n = invH.shape[0]
        order = [j for j in range(n) if j != q] + [q]                     # column / row q moved to the end (:289-291)
        P = invH[:, order][order, :]
        return P[:-1, :-1] - torch.outer(P[:-1, -1], P[-1, :-1]) / P[-1, -1]


def ecm_weights(C, tolerance=1e-8, *, device=None):
    """Weights the way run_HPROM_ecm.py:80-91 trains them: a basis ``u`` of the columns space of ``C.T``
    (left singular vectors, singular values below ``tolerance`` x the largest dropped), ECM with the sum-of-
    weights constraint, zeros elsewhere.  ``C``: (k n_snaps, n_cells) NumPy array or tensor; returns the NumPy
    weight vector (n_cells,), the selector (``z``, ``w``, ``nerror``)."""
    Ct = C if isinstance(C, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(C, dtype=np.float64))
    if device is not None:
        Ct = Ct.to(device)
    q, r = torch.linalg.qr(Ct.T.contiguous(), mode="reduced")             # thin SVD of C^T through QR (as pod.py does)
    ur, s, _ = torch.linalg.svd(r, full_matrices=False)
    keep = s > tolerance * s[0]
    u = q @ ur[:, keep]
    sel = EmpiricalCubatureMethod()
    sel.SetUp(u, constrain_sum_of_weights=True, constrain_conditions=False)
    sel.Run()
    weights = np.zeros(Ct.shape[1])
    weights[sel.z] = np.squeeze(sel.w)
    return weights, sel
