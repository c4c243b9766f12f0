"""Parity at the sizes BASELINE.json's configs name, against the UNMODIFIED reference.

Fixtures: oracle/gen_golden_large.py (reference imported as-is in the build container).
  * FOM over the full 500 steps at 64^2 and 250^2 (config 1: BurgersFD_CleanCoarse/run_fom.py:9):
    Newton counts of every step equal, every snapshot norm and the sampled entries within 1e-10;
  * LSPG at 250^2 and 750^2 with k = 95 (config 4's sizes) on the analytic hat basis, and at 250^2
    on the POD basis of two reference-sized trajectories;
  * the ECSW training matrix at 250^2, k = 95; the HPROM on weights the reference's own NNLS chain
    produced (run_HPROM_ecsw_joshua.py:55-100), index sets bit-exact.
Tolerances (BASELINE.json north_star): FOM 1e-10 per snapshot, reduced states 1e-8, counts equal.
"""
import numpy as np
import pytest

from oracle import fixtures as fx

pytestmark = pytest.mark.gpu
FOM_TOL = 1e-10
ROM_TOL = 1e-8


@pytest.fixture(scope="module")
def h():
    from finitedifference_b200 import hypernet2D
    return hypernet2D


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


def rel_cols(a, b):
    return (np.linalg.norm(a - b, axis=0) / np.linalg.norm(b, axis=0)).max()


@pytest.mark.parametrize("name", ["fom_n64_T500.npz", "fom_n250_T500.npz"])
def test_fom_full_trajectory(h, golden, name):
    """All 500 steps: the late trajectory (front leaving the domain, Newton counts changing,
    continuation passes and redone steps) against the reference."""
    g = golden(name)
    n, steps = int(g["n"]), int(g["steps"])
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    snaps, info = h.inviscid_burgers_implicit2D(gx, gy, np.ones(2 * n * n), float(g["dt"]), steps, list(g["mu"]),
                                                return_info=True)
    assert snaps.shape == (2 * n * n, steps + 1)
    assert list(info["newton_its"]) == list(g["newton_its"])
    assert n < 250 or len(set(g["newton_its"].tolist())) > 1            # at 250^2 the count does change along the way
    assert np.max(np.abs(np.linalg.norm(snaps, axis=0) / g["norms"] - 1)) < FOM_TOL
    assert rel_cols(snaps[::int(g["dof_stride"])], g["dof_sample"]) < FOM_TOL
    assert rel_cols(snaps[:, ::int(g["t_stride"])], g["t_sample"]) < FOM_TOL


def test_fom_training_parameter_norms(h, golden):
    g = golden("fom_n250_T500_train.npz")
    n, steps = int(g["n"]), int(g["steps"])
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    solver = h.inviscid_burgers_implicit2D_batch(gx, gy, np.ones(2 * n * n), float(g["dt"]), steps, [list(g["mu"])])
    its, _ = solver.diagnostics()
    assert its[0].tolist() == g["newton_its"].tolist()
    norms = solver.snapshots_device(0, layout=0).norm(dim=1).cpu().numpy()
    assert np.max(np.abs(norms / g["norms"] - 1)) < FOM_TOL


@pytest.mark.parametrize("name", ["lspg_hat_n250.npz", "lspg_hat_n750.npz"])
def test_lspg_k95_hat_basis(h, golden, name):
    """The k = 95 instantiations (three 32-mode chunks of the grid-walk J V, bfd_gram, bfd_spd_solve,
    the V y ring) on the coarse and the fine grid against the reference."""
    g = golden(name)
    n, steps, k = int(g["n"]), int(g["steps"]), int(g["k"])
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    V = fx.hat_basis(n, k)
    stride = int(g["stride"])
    assert np.max(np.abs(V[::stride] - g["basis_sample"])) < 1e-13     # the same basis as in the build container
    snaps, times, info = h.inviscid_burgers_implicit2D_LSPG(gx, gy, np.ones(2 * n * n), float(g["dt"]), steps,
                                                            list(g["mu"]), V, return_info=True)
    assert info["gn_its"].tolist() == g["gn_its"].tolist() and times[0] == int(g["num_its"])
    assert max(rel(info["red_coords"][:, i], g["red_coords"][:, i]) for i in range(steps + 1)) < ROM_TOL
    assert rel_cols(snaps[::stride], g["sample"]) < ROM_TOL
    assert np.max(np.abs(np.linalg.norm(snaps, axis=0) / g["norms"] - 1)) < ROM_TOL


@pytest.fixture(scope="module")
def rom250(h, golden):
    """The POD basis of the 250^2 fixture, rebuilt from W/sigma of the reference's SVD and the two
    training trajectories recomputed here (the FOM is pinned by test_fom_full_trajectory)."""
    g = golden("rom_n250_k95.npz")
    n = int(g["n"])
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    solver = h.inviscid_burgers_implicit2D_batch(gx, gy, w0, float(g["dt"]), 500, [list(m) for m in g["mu_train"]])
    trajs = [solver.snapshots(m) for m in range(2)]
    V = fx.pod_from_coef(fx.pod_training_matrix(trajs), g["coef"])
    return g, gx, gy, w0, V, trajs


def test_pod_basis_rebuilt(rom250):
    g, gx, gy, w0, V, _ = rom250
    bs = int(g["basis_stride"])
    assert np.max(np.abs(V[::bs] - g["basis_sample"])) < 1e-9
    assert np.max(np.abs(V.T @ V - np.eye(V.shape[1]))) < 1e-8


def test_lspg_k95_pod_basis(h, rom250):
    g, gx, gy, w0, V, _ = rom250
    steps = int(g["lspg_steps"])
    snaps, times, info = h.inviscid_burgers_implicit2D_LSPG(gx, gy, w0, float(g["dt"]), steps, list(g["mu"]), V,
                                                            return_info=True)
    assert info["gn_its"].tolist() == g["lspg_gn_its"].tolist() and times[0] == int(g["lspg_num_its"])
    assert rel_cols(snaps[::int(g["lspg_stride"])], g["lspg_sample"]) < ROM_TOL
    assert np.max(np.abs(np.linalg.norm(snaps, axis=0) / g["lspg_norms"] - 1)) < ROM_TOL
    assert max(rel(V.T @ snaps[:, i], g["lspg_red"][:, i]) for i in range(steps + 1)) < ROM_TOL


def test_training_matrix_k95(h, rom250):
    """compute_ECSW_training_matrix_2D at 250^2, k = 95 on the snapshots run_HPROM_ecsw_joshua.py:58-59
    selects: two blocks entry by entry, all of C through its row and column sums."""
    g, gx, gy, w0, V, trajs = rom250
    Sb, T, k = trajs[1], 500, int(g["k"])
    C = h.compute_ECSW_training_matrix_2D(Sb[:, 3:T:10], Sb[:, 0:T - 3:10], V, h.inviscid_burgers_res2D,
                                          h.inviscid_burgers_exact_jac2D, gx, gy, float(g["dt"]), list(g["mu_train"][1]))
    assert C.shape == (50 * k, 250 * 250)
    scale = np.abs(g["C_block7"]).max()
    assert np.max(np.abs(C[7 * k:8 * k, ::int(g["C_stride"])] - g["C_block7"])) < ROM_TOL * scale
    assert np.max(np.abs(C[31 * k:32 * k, 30000:30600] - g["C_block31"])) < ROM_TOL * np.abs(g["C_block31"]).max()
    assert np.max(np.abs(C.sum(axis=0) - g["C_colsum"])) < ROM_TOL * np.abs(g["C_colsum"]).max()
    assert np.max(np.abs(C.sum(axis=1) - g["C_rowsum"])) < ROM_TOL * np.abs(g["C_rowsum"]).max()


def test_ecsw_reference_weights(h, rom250):
    """HPROM on the weights the reference's own chain produced (interior mask, 20 NNLS blocks, bc_w = 10)."""
    g, gx, gy, w0, V, _ = rom250
    steps = int(g["ecsw_steps"])
    red, times, info = h.inviscid_burgers_ecsw_fixed(gx, gy, g["weights"], w0, float(g["dt"]), steps, list(g["mu"]), V,
                                                     return_info=True)
    assert np.array_equal(info["sample_inds"], g["sample_inds"])
    assert np.array_equal(info["augmented_sample"], g["augmented_sample"])
    assert info["gn_its"].tolist() == g["ecsw_gn_its"].tolist()
    assert max(rel(red[:, i], g["ecsw_red"][:, i]) for i in range(steps + 1)) < ROM_TOL
