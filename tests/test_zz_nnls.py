"""Lawson-Hanson NNLS on torch tensors (finitedifference_b200/nnls.py): host logic on CPU tensors
against scipy.optimize.nnls -- the call the reference makes (run_HPROM_ecsw_joshua.py:83) -- where the
solution is unique, KKT conditions where it is not (the reference's shape: more columns than rows)."""
import numpy as np
import pytest
import torch
from scipy.optimize import nnls as scipy_nnls

from finitedifference_b200 import nnls as N


def kkt(A, b, x, tol):
    w = A.T @ (b - A @ x)
    assert (x >= 0).all()
    assert (w[x == 0] <= tol).all()                       # no inactive variable wants to grow
    assert np.abs(w[x > 0]).max(initial=0.0) <= tol       # stationarity on the support


@pytest.mark.parametrize("m,n,seed", [(60, 25, 0), (200, 40, 1), (30, 30, 2), (50, 8, 3)])
def test_unique_solution_equals_scipy(m, n, seed):
    rng = np.random.default_rng(seed)
    A = rng.standard_normal((m, n))
    b = A @ np.where(rng.random(n) < 0.5, rng.random(n), 0.0) + 0.1 * rng.standard_normal(m)
    x, rnorm, its = N.nnls_tensor(torch.from_numpy(A), torch.from_numpy(b))
    xs, rs = scipy_nnls(A, b)
    assert np.array_equal(x.numpy() > 0, xs > 0)          # same support
    assert np.allclose(x.numpy(), xs, rtol=1e-9, atol=1e-11) and abs(rnorm - rs) <= 1e-9 * max(rs, 1.0)
    kkt(A, b, x.numpy(), 1e-9 * np.abs(A).sum())


@pytest.mark.parametrize("m,n,seed", [(20, 90, 4), (35, 400, 5)])
def test_reference_shaped_problem_kkt(m, n, seed):
    """b = A @ 1 with n > m, A >= 0 rows mixed in sign like the training matrix: a sparse positive
    solution reproduces b; which one is rounding-sensitive, so check the optimality conditions."""
    rng = np.random.default_rng(seed)
    A = rng.standard_normal((m, n)) * rng.random(n)
    b = A.sum(axis=1)
    x, rnorm, its = N.nnls_tensor(torch.from_numpy(A), torch.from_numpy(b))
    x = x.numpy()
    assert rnorm <= 1e-9 * np.linalg.norm(b) and np.count_nonzero(x) <= m
    assert scipy_nnls(A, b)[1] <= 1e-9 * np.linalg.norm(b)
    kkt(A, b, x, 1e-8 * np.abs(A).sum())


def rank_deficient_problem(seed=11, m=25, n=120):
    """Duplicated and linearly combined columns, more columns than rows, exact solution x = 1."""
    rng = np.random.default_rng(seed)
    base = rng.standard_normal((m, n // 3)) * rng.random(n // 3)
    A = np.hstack([base, base, 0.5 * base + 0.5 * np.roll(base, 1, axis=1)])
    return A, A.sum(axis=1)


def test_rank_deficient_blocks_terminate_with_kkt():
    """Dependent passive columns (the Cholesky of the passive Gram matrix breaks down): the
    minimum-norm branch, no cycling, optimality conditions hold."""
    A, b = rank_deficient_problem()
    x, rnorm, its = N.nnls_tensor(torch.from_numpy(A), torch.from_numpy(b), maxiter=9999999999)
    assert rnorm <= 1e-8 * np.linalg.norm(b) and its < 10 * A.shape[1]
    kkt(A, b, x.numpy(), 1e-7 * np.abs(A).sum())


def test_edge_cases():
    A = np.array([[1.0, 0.0], [1.0, 0.0], [0.0, 1.0]])
    x, r, _ = N.nnls_tensor(torch.from_numpy(A), torch.tensor([2.0, 1.0, 1.0], dtype=torch.float64))
    assert np.allclose(x.numpy(), [1.5, 1.0]) and abs(r - 0.7071067811865475) < 1e-14     # scipy's docstring example
    x, r, _ = N.nnls_tensor(torch.from_numpy(A), torch.tensor([-1.0, -1.0, -1.0], dtype=torch.float64))
    assert np.array_equal(x.numpy(), [0.0, 0.0]) and abs(r - 1.7320508075688772) < 1e-14
    with pytest.raises(ValueError):
        N.nnls_tensor(torch.zeros(3, 2), torch.zeros(4))
    if not torch.cuda.is_available():
        from finitedifference_b200._lib import BurgersError
        with pytest.raises(BurgersError):
            N.nnls(A, np.ones(3))


@pytest.mark.gpu
def test_device_path_and_block_training():
    rng = np.random.default_rng(7)
    A = rng.standard_normal((80, 30))
    b = A @ np.where(rng.random(30) < 0.5, rng.random(30), 0.0) + 0.05 * rng.standard_normal(80)
    x, r = N.nnls(A, b)
    xs, rs = scipy_nnls(A, b)
    assert np.allclose(x, xs, rtol=1e-8, atol=1e-10) and abs(r - rs) < 1e-9
    Ad, bd = rank_deficient_problem()                      # dependent columns on the device path
    xd, rd = N.nnls(Ad, bd, maxiter=9999999999)
    assert rd <= 1e-8 * np.linalg.norm(bd)
    kkt(Ad, bd, xd, 1e-7 * np.abs(Ad).sum())
    C = rng.standard_normal((24, 300)) * rng.random(300)
    wts = N.ecsw_weights(C, blocks=5)
    assert wts.shape == (300,) and (wts >= 0).all()
    for c, wblk in zip(np.array_split(C, 5, axis=1), np.array_split(wts, 5)):
        assert np.linalg.norm(c @ wblk - c.sum(axis=1)) <= 1e-8 * np.linalg.norm(c.sum(axis=1))
