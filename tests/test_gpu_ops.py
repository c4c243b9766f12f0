"""The operator-level callables of the drop-in module (finitedifference_b200/operators.py) on the
GPU, against golden vectors generated from the unmodified reference and against the oracle."""
import numpy as np
import pytest

from oracle import burgers_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def h():
    from finitedifference_b200 import hypernet2D
    return hypernet2D


def test_residuals_and_jacobian_golden(h, golden):
    g = golden("ops_n12.npz")
    n, dt, mu = int(g["n"]), float(g["dt"]), list(g["mu"])
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    Dxec, Dyec, JDxec, JDyec, Eye = h.get_ops(gx, gy)
    w, wp = g["w"].copy(), g["wp"].copy()
    assert np.max(np.abs(h.inviscid_burgers_res2D(w, gx, gy, dt, wp, mu, Dxec, Dyec) - g["r_std"])) < 1e-14
    assert np.max(np.abs(h.inviscid_burgers_res2D_alt(w, gx, gy, dt, wp, mu, JDxec, JDyec) - g["r_alt"])) < 1e-14
    J = h.inviscid_burgers_exact_jac2D(w, dt, JDxec, JDyec, Eye)
    assert J.format == "csr" and np.max(np.abs(J.toarray() - g["J_exact"])) < 1e-14
    qx, qy = orc.lifted_operators(gx, gy, "unpermuted")                            # hypernet2D.py:167
    assert np.max(np.abs(h.inviscid_burgers_exact_jac2D(w, dt, qx, qy, Eye).toarray() - g["J_lspg"])) < 1e-14
    assert np.array_equal(w, g["w"]) and np.array_equal(wp, g["wp"])               # inputs untouched


def test_sampled_residual_and_jacobian_golden(h, golden):
    g = golden("ecsw_ops_n24.npz")
    n, dt, mu = int(g["n"]), float(g["dt"]), list(g["mu"])
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    sample = np.where(g["weights"] != 0)[0]
    aug = h.generate_augmented_mesh(gx, gy, sample)
    jdx_s, jdy_s, eye2 = orc.ecsw_operators(gx, gy, sample, aug)
    idx = np.concatenate((aug, n * n + aug))
    w, wp = g["w"], g["wp"]
    full = h.inviscid_burgers_res2D_ecsw(w, gx, gy, dt, wp, mu, jdx_s, jdy_s, sample, aug)
    red = h.inviscid_burgers_res2D_ecsw(w[idx], gx, gy, dt, wp[idx], mu, jdx_s, jdy_s, sample, aug)
    assert np.max(np.abs(full - g["r_full"])) < 1e-14 and np.max(np.abs(red - g["r_red"])) < 1e-14
    # precomputed inlet / source terms, as the manifold drivers pass them (hypernet2D.py:630-652): the reference's
    # own formulas give the same residual; other values shift it by exactly their difference
    dx, xc = gx[1:] - gx[:-1], (gx[1:] + gx[:-1]) / 2
    lbc = np.where(sample % n == 0, 0.5 * dt * mu[0] ** 2 / dx[0], 0.0)
    src = np.tile(dt * 0.02 * np.exp(mu[1] * xc), n)[sample]
    same = h.inviscid_burgers_res2D_ecsw(w[idx], gx, gy, dt, wp[idx], mu, jdx_s, jdy_s, sample, aug, lbc, src)
    assert np.max(np.abs(same - g["r_red"])) < 1e-14
    other = h.inviscid_burgers_res2D_ecsw(w[idx], gx, gy, dt, wp[idx], mu, jdx_s, jdy_s, sample, aug, lbc + 0.25, 2 * src)
    assert np.max(np.abs(other[:sample.size] - (g["r_red"][:sample.size] - 0.25 - src))) < 1e-13
    assert np.max(np.abs(other[sample.size:] - g["r_red"][sample.size:])) < 1e-14
    for state in (w, w[idx]):
        J = h.inviscid_burgers_exact_jac2D_ecsw(state, dt, jdx_s, jdy_s, eye2, sample, aug)
        assert J.shape == g["J_red"].shape and np.max(np.abs(J.toarray() - g["J_red"])) < 1e-14


def test_jacobian_at_fine_grid_against_oracle(h):
    n, dt = 750, 0.05
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    _, _, JDxec, JDyec, Eye = h.get_ops(gx, gy)
    rng = np.random.default_rng(1)
    w = 1.0 + rng.random(2 * n * n)
    J = h.inviscid_burgers_exact_jac2D(w, dt, JDxec, JDyec, Eye)
    Jo = orc.jacobian(w, dt, *orc.lifted_operators(gx, gy))
    assert J.shape == Jo.shape and abs(J - Jo).max() < 1e-13 and J.nnz == Jo.nnz
