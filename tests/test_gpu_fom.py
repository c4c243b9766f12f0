"""Parity of the CUDA FOM path (through the drop-in API and the C ABI) with the oracle and with
the golden vectors generated from the unmodified reference.  Tolerance: per-snapshot relative
L2 error <= 1e-10 (BASELINE.json north_star); Newton iteration counts must be equal."""
import numpy as np
import pytest

from oracle import burgers_oracle as orc

pytestmark = pytest.mark.gpu
TOL = 1e-10


def rel_err(a, b):
    return (np.linalg.norm(a - b, axis=0) / np.linalg.norm(b, axis=0)).max()


@pytest.fixture(scope="module")
def h():
    from finitedifference_b200 import hypernet2D
    return hypernet2D


@pytest.mark.parametrize("n", [16, 24, 48])
@pytest.mark.parametrize("kw", [{}, dict(chunk_steps=3, newton_hint=6), dict(chunk_steps=1, newton_hint=1)])
def test_fom_golden(h, golden, n, kw):
    g = golden(f"fom_n{n}.npz")
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    snaps, info = h.inviscid_burgers_implicit2D(gx, gy, w0, float(g["dt"]), int(g["steps"]), list(g["mu"]),
                                                return_info=True, **kw)
    assert snaps.shape == g["snaps"].shape and snaps.dtype == np.float64 and snaps.flags.c_contiguous
    assert list(info["newton_its"]) == list(g["newton_its"])
    assert rel_err(snaps, g["snaps"]) < TOL
    assert np.array_equal(w0, np.ones(2 * n * n))          # inputs are not mutated


@pytest.mark.parametrize("nx,ny", [(100, 100), (130, 45), (33, 97), (1, 1), (5, 1), (1, 7), (32, 8), (65, 9)])
def test_fom_vs_oracle(h, nx, ny):
    gx, gy = h.make_2D_grid(0, 100, 0, 100, nx, ny)
    rng = np.random.default_rng(nx * 1000 + ny)
    w0 = 1.0 + 0.5 * rng.random(2 * nx * ny)               # non-trivial initial state
    mu = [5.1, 0.024]
    snaps, info = h.inviscid_burgers_implicit2D(gx, gy, w0, 0.05, 3, mu, return_info=True)
    ref, hist = orc.fom_trajectory(gx, gy, w0, 0.05, 3, mu, return_info=True)
    assert list(info["newton_its"]) == [len(x) for x in hist]
    assert rel_err(snaps, ref) < TOL


def test_fom_batch_of_mu(h):
    n = 64
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    mus = [[m1, m2] for m1 in np.linspace(4.25, 5.5, 3) for m2 in np.linspace(0.015, 0.03, 3)]
    w0 = np.ones(2 * n * n)
    solver = h.inviscid_burgers_implicit2D_batch(gx, gy, w0, 0.05, 6, mus, chunk_steps=4)
    its, rel = solver.diagnostics()
    for m in (0, 4, 8):
        ref, hist = orc.fom_trajectory(gx, gy, w0, 0.05, 6, mus[m], return_info=True)
        assert its[m].tolist() == [len(x) for x in hist]
        assert rel_err(solver.snapshots(m), ref) < TOL
        dev = solver.snapshots_device(m, t0=2, t1=5, layout=0).cpu().numpy()
        assert np.array_equal(dev, solver.snapshots(m)[:, 2:5].T)
    assert np.all(rel < 1e-12)


def test_zero_steps_and_w0_per_sample(h):
    n = 20
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    snaps = h.inviscid_burgers_implicit2D(gx, gy, w0, 0.05, 0, [4.75, 0.02])
    assert snaps.shape == (2 * n * n, 1) and np.array_equal(snaps[:, 0], w0)
    from finitedifference_b200.fom import FomSolver
    rng = np.random.default_rng(1)
    w0s = 1 + rng.random((2, 2 * n * n))
    s = FomSolver(gx, gy, 0.05, 2, [[4.75, 0.02], [4.75, 0.02]])
    s.set_initial(w0s)
    s.run()
    for m in range(2):
        ref = orc.fom_trajectory(gx, gy, w0s[m], 0.05, 2, [4.75, 0.02])
        assert rel_err(s.snapshots(m), ref) < TOL


@pytest.mark.parametrize("n", [250, 750])
def test_fom_reference_grids_first_steps(h, golden, n):
    """First steps of the coarse (250^2) and fine (750^2) configs against the reference itself."""
    g = golden(f"fom_n{n}.npz")
    steps, stride = int(g["steps"]), int(g["stride"])
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    snaps, info = h.inviscid_burgers_implicit2D(gx, gy, np.ones(2 * n * n), float(g["dt"]), steps, list(g["mu"]),
                                                return_info=True)
    assert list(info["newton_its"]) == list(g["newton_its"])
    assert np.max(np.abs(np.linalg.norm(snaps, axis=0) / g["norms"] - 1)) < TOL
    assert rel_err(snaps[::stride], g["sample"]) < TOL


def test_fine_grid_snapshots_solve_their_step(h):
    """Size-independent property at the full fine size: every snapshot is a root of its step's
    residual to the Newton tolerance, checked with the natural-layout residual kernel."""
    import ctypes as C
    import torch
    from finitedifference_b200 import _lib
    from finitedifference_b200.discretisation import parameter_terms, stencil_coefficients
    n, steps, mu = 750, 4, [4.75, 0.02]
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    solver = h.inviscid_burgers_implicit2D_batch(gx, gy, np.ones(2 * n * n), 0.05, steps, [mu, [5.5, 0.03]])
    lib = _lib.get_lib()
    dev = solver.device
    coefs = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in stencil_coefficients(gx, gy, 0.05)]
    for m, mm in enumerate([mu, [5.5, 0.03]]):
        src, lbc = (torch.from_numpy(a).to(dev) for a in parameter_terms(gx, gy, 0.05, [mm]))
        S = solver.snapshots_device(m, layout=0)
        r = torch.empty(2 * n * n, dtype=torch.float64, device=dev)
        for i in range(steps):
            norms = []
            for w in (S[i], S[i + 1]):
                n2 = torch.zeros(1, dtype=torch.float64, device=dev)
                rc = lib.bfd_residual(n, n, *[c.data_ptr() for c in coefs], src.data_ptr(), lbc.data_ptr(),
                                      w.data_ptr(), S[i].data_ptr(), r.data_ptr(), n2.data_ptr(),
                                      torch.cuda.current_stream().cuda_stream)
                assert rc == 0
                norms.append(float(n2.sqrt()))
            assert norms[1] / norms[0] < 1e-12


def test_snapshot_cache_is_reference_compatible(h, tmp_path):
    n = 24
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    folder = str(tmp_path / "param_snaps")
    a = h.load_or_compute_snaps([4.25, 0.015], gx, gy, w0, 0.05, 4, snap_folder=folder)
    import os
    assert os.path.exists(os.path.join(folder, "mu1_4.25+mu2_0.015.npy"))
    b = h.load_or_compute_snaps([4.25, 0.015], gx, gy, w0, 0.05, 3, snap_folder=folder)
    assert np.array_equal(a[:, :4], b)
    on_disk = np.load(os.path.join(folder, "mu1_4.25+mu2_0.015.npy"))
    assert on_disk.shape == (2 * n * n, 5) and on_disk.flags.c_contiguous


@pytest.mark.parametrize("n,steps,kw", [(64, 20, {}), (100, 7, dict(chunk_steps=2, newton_hint=1)), (250, 12, {})])
def test_tile_kernel_equals_chunk_kernel(h, n, steps, kw):
    """fom_tile_kernel (one warp per strip and time step, bfd_fom_desc.kernel = 1) against the default
    fom_chunk_kernel: same Newton counts, same snapshots (the arithmetic differs only by exact
    power-of-two rescalings); at most 4 solves are fused per pass, further ones are continuation passes."""
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    a, ia = h.inviscid_burgers_implicit2D(gx, gy, w0, 0.05, steps, [5.19, 0.026], return_info=True, **kw)
    b, ib = h.inviscid_burgers_implicit2D(gx, gy, w0, 0.05, steps, [5.19, 0.026], return_info=True, kernel="tile", **kw)
    assert list(ia["newton_its"]) == list(ib["newton_its"])
    assert rel_err(b, a) < 1e-14
