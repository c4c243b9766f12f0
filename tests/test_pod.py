"""POD drop-in (finitedifference_b200/pod.py): host logic on CPU tensors against NumPy's SVD and
against scikit-learn's randomized_svd -- the two calls the reference's POD makes
(hypernet2D.py:2746-2771) -- and the device path on the GPU."""
import numpy as np
import pytest
import torch

from finitedifference_b200 import pod


def snapshot_like(m, n, seed=0):
    """Decaying-spectrum matrix, like FOM snapshot sets."""
    rng = np.random.default_rng(seed)
    u, _ = np.linalg.qr(rng.standard_normal((m, n)))
    v, _ = np.linalg.qr(rng.standard_normal((n, n)))
    s = 10.0 ** (-np.arange(n) / 4.0) * 1e3
    return (u * s) @ v.T


def check_svd(u, s, A, k):
    ur, sr, _ = np.linalg.svd(A, full_matrices=False)
    assert np.allclose(s, sr, rtol=1e-10, atol=1e-12 * sr[0])
    overlap = np.abs(np.sum(u[:, :k] * ur[:, :k], axis=0))          # same vectors up to sign
    assert np.all(overlap > 1 - 1e-8), overlap.min()
    assert np.abs(u.T @ u - np.eye(u.shape[1])).max() < 1e-12


def check_rsvd(u, s, A, k, seed):
    from sklearn.utils.extmath import randomized_svd
    ur, sr, _ = randomized_svd(A, n_components=k, random_state=seed)
    assert u.shape == ur.shape and np.allclose(s, sr, rtol=1e-9)
    good = sr > 1e-6 * sr[0]                                         # well-separated part of the spectrum
    assert np.max(np.abs(u[:, good] - ur[:, good])) < 1e-6           # same vectors, same signs (svd_flip)


@pytest.mark.parametrize("shape", [(400, 30), (30, 400)])
def test_pod_host_logic(shape):
    A = snapshot_like(*shape) if shape[0] > shape[1] else snapshot_like(*shape[::-1]).T
    u, s = pod.pod_tensor(torch.from_numpy(A))
    check_svd(u.numpy(), s.numpy(), A, 12)
    for k, seed in ((5, 0), (2, 7)):                                 # n_iter 4 and 7 of 'auto'
        u, s = pod.pod_tensor(torch.from_numpy(A), k, "rsvd", seed)
        check_rsvd(u.numpy(), s.numpy(), A, k, seed)
    with pytest.raises(ValueError):
        pod.pod_tensor(torch.from_numpy(A), method="qr")


def test_pod_refuses_cpu():
    from finitedifference_b200._lib import BurgersError
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    with pytest.raises(BurgersError):
        pod.POD(np.ones((8, 3)))


@pytest.mark.gpu
def test_pod_device():
    A = snapshot_like(20000, 60, seed=3)
    u, s = pod.POD(A)
    assert isinstance(u, np.ndarray) and u.shape == (20000, 60)
    check_svd(u, s, A, 25)
    u, s = pod.POD(A, 8, method="rsvd", random_state=0)
    check_rsvd(u, s, A, 8, 0)
    ud, sd = pod.POD(torch.from_numpy(A).cuda(), 8, method="rsvd", random_state=0, out_device=True)
    assert ud.is_cuda and np.allclose(ud.cpu().numpy(), u, atol=1e-9)


@pytest.mark.gpu
def test_pod_of_fom_snapshots_feeds_lspg():
    """run_prom.py's chain on the device: FOM snapshots -> POD -> LSPG; the PROM reproduces the
    FOM it was trained on to the projection error."""
    from finitedifference_b200 import hypernet2D as h
    n, steps, mu = 32, 10, [4.75, 0.02]
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    solver = h.inviscid_burgers_implicit2D_batch(gx, gy, w0, 0.05, steps, [mu])
    S = solver.snapshots_device(0, layout=1)                          # (2N, T+1) in HBM
    u, s = h.POD(S, out_device=True)
    assert float((u.T @ u - torch.eye(u.shape[1], device=u.device, dtype=u.dtype)).abs().max()) < 1e-12
    basis = u[:, :8].contiguous()
    prom, _ = h.inviscid_burgers_implicit2D_LSPG(gx, gy, w0, 0.05, steps, mu, basis.cpu().numpy())
    fom = S.cpu().numpy()
    rel, mean = h.compute_error(prom, fom)
    assert mean < 1e-3
