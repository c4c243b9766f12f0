"""Pin the CPU oracle (oracle/burgers_oracle.py) against outputs of the unmodified reference.

The fixtures under tests/golden/ were produced by oracle/gen_golden.py, which imports
/root/reference/BurgersFD_CleanFine/hypernet2D.py as-is.  CPU only.
"""
import numpy as np
import pytest

from oracle import burgers_oracle as orc


def rel_l2_cols(a, b):
    return np.linalg.norm(a - b, axis=0) / np.linalg.norm(b, axis=0)


def grid(n):
    return orc.make_2D_grid(0, 100, 0, 100, n, n)


def test_residual_and_jacobians(golden):
    g = golden("ops_n12.npz")
    n, dt, mu = int(g["n"]), float(g["dt"]), list(g["mu"])
    gx, gy = grid(n)
    r = orc.residual(g["w"], gx, gy, dt, g["wp"], mu)
    # res2D and res2D_alt of the reference agree to rounding; the oracle follows _alt
    assert np.max(np.abs(g["r_alt"] - g["r_std"])) < 1e-14
    assert np.max(np.abs(r - g["r_alt"])) < 1e-15
    jdx, jdy = orc.lifted_operators(gx, gy)
    J = orc.jacobian(g["w"], dt, jdx, jdy).toarray()
    assert np.max(np.abs(J - g["J_exact"])) < 1e-15
    jdxq, jdyq = orc.lifted_operators(gx, gy, "unpermuted")
    Jq = orc.jacobian(g["w"], dt, jdxq, jdyq).toarray()
    assert np.max(np.abs(Jq - g["J_lspg"])) < 1e-15
    # the FOM Jacobian is 2x2-block lower triangular in cell-interleaved ordering
    perm = np.arange(2 * n * n).reshape(2, -1).T.ravel()
    Jp = J[np.ix_(perm, perm)]
    blocks_above = np.triu(np.kron(np.ones((n * n, n * n)), np.ones((2, 2))), 2)
    assert np.count_nonzero(Jp * (np.triu(np.ones_like(Jp), 2) * (blocks_above > 0))) == 0


@pytest.mark.parametrize("n", [16, 24, 48])
def test_fom_trajectory(golden, n):
    g = golden(f"fom_n{n}.npz")
    gx, gy = grid(n)
    snaps, hist = orc.fom_trajectory(gx, gy, np.ones(2 * n * n), float(g["dt"]), int(g["steps"]),
                                     list(g["mu"]), return_info=True)
    assert [len(h) for h in hist] == list(g["newton_its"])
    assert rel_l2_cols(snaps, g["snaps"]).max() < 1e-13


def test_lspg_trajectory(golden):
    g = golden("lspg_n24.npz")
    n = int(g["n"])
    gx, gy = grid(n)
    snaps, times, red, hist = orc.lspg_trajectory(gx, gy, np.ones(2 * n * n), float(g["dt"]),
                                                  int(g["steps"]), list(g["mu"]), g["basis"],
                                                  return_info=True)
    assert [len(h) for h in hist] == list(g["gn_its"])
    assert times[0] == int(g["num_its"])
    assert rel_l2_cols(snaps, g["snaps"]).max() < 1e-12
    # the exact-Jacobian switch is a different (mathematically right) method: results differ
    snaps_e, _ = orc.lspg_trajectory(gx, gy, np.ones(2 * n * n), float(g["dt"]), int(g["steps"]),
                                     list(g["mu"]), g["basis"], jacobian_mode="exact")
    assert rel_l2_cols(snaps_e, g["snaps"]).max() > 1e-9


def test_ecsw_index_sets_and_trajectory(golden):
    g = golden("ecsw_n24.npz")
    n = int(g["n"])
    gx, gy = grid(n)
    red, times, sample_inds, aug, hist = orc.ecsw_trajectory(
        gx, gy, g["weights"], np.ones(2 * n * n), float(g["dt"]), int(g["steps"]), list(g["mu"]),
        g["basis"], return_info=True)
    assert sample_inds.dtype == np.int64 and np.array_equal(sample_inds, g["sample_inds"])
    assert aug.dtype == np.int64 and np.array_equal(aug, g["augmented_sample"])
    assert [len(h) for h in hist] == list(g["gn_its"])
    assert rel_l2_cols(red, g["red_coords"]).max() < 1e-11


def test_ecsw_operators(golden):
    g = golden("ecsw_ops_n24.npz")
    n, dt, mu = int(g["n"]), float(g["dt"]), list(g["mu"])
    gx, gy = grid(n)
    s = np.where(g["weights"] != 0)[0]
    aug = orc.augmented_mesh(gx, gy, s)
    jdx_s, jdy_s, eye2 = orc.ecsw_operators(gx, gy, s, aug)
    idx = np.concatenate((aug, n * n + aug))
    w, wp = g["w"], g["wp"]
    r_full = orc.residual_ecsw(w, gx, gy, dt, wp, mu, jdx_s, jdy_s, s, aug)
    r_red = orc.residual_ecsw(w[idx], gx, gy, dt, wp[idx], mu, jdx_s, jdy_s, s, aug)
    assert np.max(np.abs(r_full - g["r_full"])) < 1e-15
    assert np.max(np.abs(r_red - g["r_red"])) < 1e-15
    # sampled residual = rows [s, N+s] of the full one (identity 3)
    rf = orc.residual(w, gx, gy, dt, wp, mu)
    assert np.max(np.abs(r_red - rf[np.concatenate((s, n * n + s))])) < 1e-14
    J = orc.jacobian_ecsw(w[idx], dt, jdx_s, jdy_s, eye2, s, aug).toarray()
    assert np.max(np.abs(J - g["J_red"])) < 1e-15


def test_ecsw_training_matrix(golden):
    g = golden("ecswC_n12.npz")
    n, steps = int(g["n"]), int(g["steps"])
    gx, gy = grid(n)
    S = g["snaps"]
    C = orc.ecsw_training_matrix(S[:, 3:steps:2], S[:, 0:steps - 3:2], g["basis"], gx, gy,
                                 float(g["dt"]), list(g["mu"]))
    assert C.shape == g["C"].shape
    assert np.max(np.abs(C - g["C"])) <= 1e-15 * max(1.0, np.max(np.abs(g["C"])))


def test_jacobian_is_derivative_of_residual():
    n, dt, mu = 10, 0.05, [4.6, 0.02]
    gx, gy = grid(n)
    rng = np.random.default_rng(3)
    w, wp = 1 + rng.random(2 * n * n), 1 + rng.random(2 * n * n)
    jdx, jdy = orc.lifted_operators(gx, gy)
    J = orc.jacobian(w, dt, jdx, jdy).toarray()
    eps = 1e-6
    fd = np.empty_like(J)
    for j in range(w.size):
        e = np.zeros_like(w); e[j] = eps
        fd[:, j] = (orc.residual(w + e, gx, gy, dt, wp, mu) - orc.residual(w - e, gx, gy, dt, wp, mu)) / (2 * eps)
    assert np.max(np.abs(J - fd)) < 1e-8


def test_snapshot_file_name():
    assert orc.param_to_snap_fn([4.25, 0.015]) == "param_snaps/mu1_4.25+mu2_0.015.npy"


def test_fom_long_trajectory_first_steps(golden):
    """The 500-step fixture of the coarse config's parameters at 64^2: the oracle over its first 60 steps."""
    g = golden("fom_n64_T500.npz")
    n, T = int(g["n"]), 60
    gx, gy = grid(n)
    snaps, hist = orc.fom_trajectory(gx, gy, np.ones(2 * n * n), float(g["dt"]), T, list(g["mu"]), return_info=True)
    assert [len(h) for h in hist] == list(g["newton_its"][:T])
    assert np.max(np.abs(np.linalg.norm(snaps, axis=0) / g["norms"][:T + 1] - 1)) < 1e-13
    assert rel_l2_cols(snaps[::int(g["dof_stride"])], g["dof_sample"][:, :T + 1]).max() < 1e-12
    ts = int(g["t_stride"])
    assert rel_l2_cols(snaps[:, ::ts], g["t_sample"][:, :T // ts + 1]).max() < 1e-12


def test_lspg_k95_at_the_coarse_grid(golden):
    """LSPG with k = 95 at 250^2 (analytic hat basis, oracle/fixtures.py) against the reference."""
    from oracle import fixtures as fx
    g = golden("lspg_hat_n250.npz")
    n, steps = int(g["n"]), 2
    gx, gy = grid(n)
    V = fx.hat_basis(n, int(g["k"]))
    assert np.max(np.abs(V[::int(g["stride"])] - g["basis_sample"])) < 1e-13
    snaps, times, red, hist = orc.lspg_trajectory(gx, gy, np.ones(2 * n * n), float(g["dt"]), steps, list(g["mu"]), V,
                                                  return_info=True)
    assert [len(h) for h in hist] == list(g["gn_its"][:steps])
    assert rel_l2_cols(red, g["red_coords"][:, :steps + 1]).max() < 1e-10
    assert rel_l2_cols(snaps[::int(g["stride"])], g["sample"][:, :steps + 1]).max() < 1e-10


def manifold_net(g):
    """The seeded decoder network of the fixture, rebuilt from the stored weights."""
    import torch
    q, nbar = g["basis"].shape[1], g["basis2"].shape[1]
    net = torch.nn.Sequential(torch.nn.Linear(q + 2, 10), torch.nn.ELU(), torch.nn.Linear(10, nbar))
    net.load_state_dict({k: torch.tensor(g["net_" + k.replace(".", "_")]) for k in net.state_dict()})
    return net


def test_manifold_hprom_trajectory(golden):
    """inviscid_burgers_rnm2D_ecsw (:556-674) restated: float32 decoder as in the reference, so the
    reduced coordinates agree to float32 rounding (1e-4 relative), Gauss-Newton counts equal."""
    g = golden("manifold_n24.npz")
    n, steps = int(g["n"]), int(g["steps"])
    gx, gy = grid(n)
    red, hist = orc.manifold_ecsw_trajectory(gx, gy, np.ones(2 * n * n), float(g["dt"]), steps, list(g["mu"]),
                                             manifold_net(g), g["basis"], g["basis2"], g["weights"])
    assert [len(h) for h in hist] == list(g["gn_its"])
    assert rel_l2_cols(red, g["red_coords"]).max() < 1e-4
