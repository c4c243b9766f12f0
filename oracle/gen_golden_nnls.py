"""Golden vectors for the two extra weight trainers (test infrastructure; run in the container that has
/root/reference):

  lsqnonneg   the UNMODIFIED reference function, imported from BurgersFD_CleanCoarse/lsqnonneg.py, on its own
              known-answer matrices (:113-172, MATLAB's resnorm values are in the file) and on two ECSW-shaped
              problems with the early stops in play (default rel_err_thresh = 0.01; max_support)
  multilevel  run_HPROM_ecsw_multilevel_complete.py:95-177 has no function to import (script body): the level
              logic is restated here line by line around scipy.optimize.nnls -- the call the script makes -- and
              its outputs stored

    python oracle/gen_golden_nnls.py        -> tests/golden/lsqnonneg_ref.npz, tests/golden/multilevel_ref.npz
"""
import contextlib
import io
import os
import sys

import numpy as np
from scipy.optimize import nnls

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/BurgersFD_CleanCoarse"


def kat_problems():
    """The matrices of the reference's own self-test (lsqnonneg.py:113-172) with MATLAB's resnorm."""
    C = np.array([[0.0372, 0.2869], [0.6861, 0.7071], [0.6233, 0.6245], [0.6344, 0.6170]])
    C1 = np.array([[0.0372, 0.2869, 0.4], [0.6861, 0.7071, 0.3], [0.6233, 0.6245, 0.1], [0.6344, 0.6170, 0.5]])
    C2 = np.array([[0.0372, 0.2869, 0.4], [0.6861, 0.7071, -0.3], [0.6233, 0.6245, -0.1], [0.6344, 0.6170, 0.5]])
    d = np.array([0.8587, 0.1781, 0.0747, 0.8405])
    k = np.array([[0.1210, 0.2319, 0.4398, 0.9342, 0.1370], [0.4508, 0.2393, 0.3400, 0.2644, 0.8188],
                  [0.7159, 0.0498, 0.3142, 0.1603, 0.4302], [0.8928, 0.0784, 0.3651, 0.8729, 0.8903],
                  [0.2731, 0.6408, 0.3932, 0.2379, 0.7349], [0.2548, 0.1909, 0.5915, 0.6458, 0.6873],
                  [0.8656, 0.8439, 0.1197, 0.9669, 0.3461], [0.2324, 0.1739, 0.0381, 0.6649, 0.1660],
                  [0.8049, 0.1708, 0.4586, 0.8704, 0.1556], [0.9084, 0.9943, 0.8699, 0.0099, 0.1911]])
    l = np.array([0.4225, 0.8560, 0.4902, 0.8159, 0.4608, 0.4574, 0.4507, 0.4122, 0.9016, 0.0056])
    C5 = np.array([[1.0, 1.0], [2.0, 1.0], [5.0, 1.0], [6.0, 1.0], [10.0, 1.0]])
    d5 = np.array([3.0, 5.0, 11.0, 13.0, 21.0])
    return [(C, d, 0.8315), (C1, d, 0.1477), (C2, d, 0.1027), (k, l, 0.3695), (k - 0.5, l, 2.8639), (C5, d5, np.nan)]


def ecsw_shaped(seed, m, n):
    rng = np.random.default_rng(seed)
    A = rng.standard_normal((m, n)) * rng.random(n)
    return A, A.sum(axis=1)


def multilevel_script(C_interior, n1, n2):
    """run_HPROM_ecsw_multilevel_complete.py:95-172, the script's own statements (joblib replaced by a loop)."""
    def solve(C_i, start_idx):
        w_i, _ = nnls(C_i, C_i.sum(axis=1), maxiter=9999999999)
        nz = np.nonzero(w_i)[0]
        return w_i[nz], np.arange(start_idx, start_idx + C_i.shape[1])[nz]
    sub1 = np.array_split(C_interior, n1, axis=1)
    st1 = np.cumsum([0] + [c.shape[1] for c in sub1[:-1]])
    r1 = [solve(c, s) for c, s in zip(sub1, st1)]
    idx1 = np.concatenate([r[1] for r in r1])
    C2 = C_interior[:, idx1]
    sub2 = np.array_split(C2, n2, axis=1)
    st2 = np.cumsum([0] + [c.shape[1] for c in sub2[:-1]])
    r2 = [solve(c, s) for c, s in zip(sub2, st2)]
    idx2 = np.concatenate([r[1] for r in r2])
    idx3 = idx1[idx2]
    z3 = np.concatenate([r[0] for r in r2])
    C3 = C_interior[:, idx3]
    b3 = C3 @ z3
    w3, res3 = nnls(C3, b3, maxiter=9999999999)
    w = np.zeros(C_interior.shape[1])
    w[idx3] = w3
    return w, res3, np.array([idx1.size, idx3.size])


def main():
    sys.path.insert(0, REF)
    from lsqnonneg import lsqnonneg          # the reference's own function, unmodified
    out = {}
    sink = io.StringIO()
    with contextlib.redirect_stdout(sink):   # the reference prints every iteration
        for i, (C, d, matlab) in enumerate(kat_problems()):
            x, resnorm, resid = lsqnonneg(C.copy(), d.copy())
            out.update({f"kat{i}_C": C, f"kat{i}_d": d, f"kat{i}_x": x, f"kat{i}_resnorm": resnorm,
                        f"kat{i}_resid": resid, f"kat{i}_matlab": matlab})
        for i, (seed, m, n, kw) in enumerate([(21, 24, 150, {}), (22, 30, 200, dict(rel_err_thresh=1e-12)),
                                              (23, 30, 200, dict(rel_err_thresh=1e-12, max_support=12))]):
            C, d = ecsw_shaped(seed, m, n)
            x, resnorm, resid = lsqnonneg(C.copy(), d.copy(), **kw)
            out.update({f"ecsw{i}_C": C, f"ecsw{i}_d": d, f"ecsw{i}_x": np.array(x), f"ecsw{i}_resnorm": resnorm,
                        f"ecsw{i}_rel_err_thresh": kw.get("rel_err_thresh", 0.01),
                        f"ecsw{i}_max_support": kw.get("max_support", -1)})
        # a problem in which a variable LEAVES the passive set (the inner loop of :66-84 runs)
        rng = np.random.default_rng(71)
        C, d = rng.standard_normal((6, 5)), rng.standard_normal(6)
        x, resnorm, resid = lsqnonneg(C.copy(), d.copy(), rel_err_thresh=0.0)
        raised = False
        try:
            lsqnonneg(C.copy(), d.copy(), itmax_factor=0, rel_err_thresh=0.0)
        except Exception:                         # "Iteration limit (1) exceeded" (:68-71)
            raised = True
        out.update({"leave_C": C, "leave_d": d, "leave_x": np.array(x), "leave_resnorm": resnorm,
                    "leave_raises_with_itmax_0": raised})
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "lsqnonneg_ref.npz"), **out)
    C, _ = ecsw_shaped(31, 30, 960)
    w, res3, sizes = multilevel_script(C, 12, 2)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "multilevel_ref.npz"), C=C, w=w, res3=res3, sizes=sizes,
                        levels=np.array([12, 2]))
    print("lsqnonneg:", {k: float(v) for k, v in out.items() if k.endswith("resnorm")})
    print("multilevel: alive", sizes, "nnz", int((w > 0).sum()), "res3", res3)


if __name__ == "__main__":
    main()
