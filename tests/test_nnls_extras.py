"""The reference's two other weight trainers, restated over the NNLS pieces (finitedifference_b200/nnls.py):
``lsqnonneg`` (BurgersFD_CleanCoarse/lsqnonneg.py:4) against outputs of the UNMODIFIED reference function and the
MATLAB values its own self-test holds (:113-172), and the three-level block training of
run_HPROM_ecsw_multilevel_complete.py:95-177 against the script's statements run with SciPy
(oracle/gen_golden_nnls.py made both fixtures).  Host logic on CPU tensors; the public entry points need CUDA."""
import numpy as np
import pytest
import torch

from finitedifference_b200 import nnls as N


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64))


@pytest.mark.parametrize("i", range(6))
def test_lsqnonneg_known_answers(golden, i):
    g = golden("lsqnonneg_ref.npz")
    C, d = g[f"kat{i}_C"], g[f"kat{i}_d"]
    x, resnorm, resid = N.lsqnonneg_tensor(T(C), T(d))
    assert np.allclose(x.numpy(), g[f"kat{i}_x"], rtol=1e-9, atol=1e-12)
    assert abs(resnorm - float(g[f"kat{i}_resnorm"])) <= 1e-10 * max(1.0, float(g[f"kat{i}_resnorm"]))
    assert np.allclose(resid.numpy(), g[f"kat{i}_resid"], rtol=1e-8, atol=1e-11)
    matlab = float(g[f"kat{i}_matlab"])
    if np.isfinite(matlab):                       # the bar of the reference's own self-test (:131-167)
        assert abs(resnorm - matlab) < 1e-3


@pytest.mark.parametrize("i", range(3))
def test_lsqnonneg_early_stops_like_the_reference(golden, i):
    """ECSW-shaped problems (more columns than rows, d = C @ 1): default 1 % relative-error stop, run to the
    end, and the cap on the support.  The stops are decided per outer iteration, so the iterate at which the
    reference stopped must be reproduced."""
    g = golden("lsqnonneg_ref.npz")
    C, d = g[f"ecsw{i}_C"], g[f"ecsw{i}_d"]
    ms = int(g[f"ecsw{i}_max_support"])
    x, resnorm, resid = N.lsqnonneg_tensor(T(C), T(d), rel_err_thresh=float(g[f"ecsw{i}_rel_err_thresh"]),
                                           max_support=ms if ms > 0 else None)
    xr = g[f"ecsw{i}_x"]
    assert np.array_equal(x.numpy() > 0, xr > 0)
    assert np.allclose(x.numpy(), xr, rtol=1e-7, atol=1e-10)
    ref = float(g[f"ecsw{i}_resnorm"])
    assert abs(resnorm - ref) <= 1e-7 * max(ref, 1e-12) + 1e-20
    if ms > 0:
        assert int((x > 0).sum()) >= ms                                  # stopped by the cap ...
        assert resnorm > 1e-6 * float(d @ d)                             # ... not by convergence


def test_lsqnonneg_variable_leaves_and_iteration_limit(golden):
    """A problem whose passive set loses a variable (the reference's inner loop, :66-84): same solution, and with
    ``itmax_factor = 0`` the first inner iteration raises, as the reference does (:68-71)."""
    g = golden("lsqnonneg_ref.npz")
    C, d = T(g["leave_C"]), T(g["leave_d"])
    x, resnorm, _ = N.lsqnonneg_tensor(C, d, rel_err_thresh=0.0)
    assert np.allclose(x.numpy(), g["leave_x"], rtol=1e-9, atol=1e-12)
    assert abs(resnorm - float(g["leave_resnorm"])) <= 1e-10 * float(g["leave_resnorm"])
    assert bool(g["leave_raises_with_itmax_0"])
    with pytest.raises(RuntimeError, match="Iteration limit"):
        N.lsqnonneg_tensor(C, d, itmax_factor=0, rel_err_thresh=0.0)


def test_multilevel_training_equals_the_script(golden):
    g = golden("multilevel_ref.npz")
    C = g["C"]
    w, rnorm, sizes = N.multilevel_weights_tensor(T(C), tuple(int(v) for v in g["levels"]))
    assert sizes == [int(v) for v in g["sizes"]]                        # columns alive after level 1 and level 2
    assert np.array_equal(w.numpy() > 0, g["w"] > 0)
    assert np.allclose(w.numpy(), g["w"], rtol=1e-7, atol=1e-10)
    assert rnorm <= 1e-9 * np.linalg.norm(C.sum(axis=1))
    # NOT asserted: C @ w == C.sum(axis=1).  The script trains level 2 against the plain column sums of the
    # surviving columns (:140), not against the level-1 quadrature it forms at :131, so its final weights
    # integrate that sum; the restatement keeps the script's behaviour.


def test_multilevel_without_levels_is_plain_nnls():
    rng = np.random.default_rng(5)
    C = rng.standard_normal((40, 12))
    w, rnorm, sizes = N.multilevel_weights_tensor(T(C), ())
    x, r, _ = N.nnls_tensor(T(C), T(C.sum(axis=1)))
    assert sizes == [] and np.allclose(w.numpy(), x.numpy()) and abs(rnorm - r) < 1e-12


def test_public_entry_points_need_cuda():
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from finitedifference_b200._lib import BurgersError
    with pytest.raises(BurgersError):
        N.lsqnonneg(np.eye(2), np.ones(2))
    with pytest.raises(BurgersError):
        N.ecsw_weights_multilevel(np.ones((2, 4)))


@pytest.mark.gpu
def test_lsqnonneg_and_multilevel_on_the_device(golden):
    g = golden("lsqnonneg_ref.npz")
    for i in range(5):
        x, resnorm, resid = N.lsqnonneg(g[f"kat{i}_C"], g[f"kat{i}_d"])
        assert abs(resnorm - float(g[f"kat{i}_matlab"])) < 1e-3 and (x >= 0).all() and resid.shape == g[f"kat{i}_d"].shape
    m = golden("multilevel_ref.npz")
    w = N.ecsw_weights_multilevel(m["C"], (12, 2))
    assert (w >= 0).all() and np.count_nonzero(w) <= m["C"].shape[0]
    if np.array_equal(w > 0, m["w"] > 0):          # same active sets on the device (rounding-sensitive): same weights
        assert np.allclose(w, m["w"], rtol=1e-6, atol=1e-9)
