"""Generate the golden fixtures under tests/golden/ by running the UNMODIFIED reference.

Runs only in the build container (needs /root/reference).  The reference's
``hypernet2D.py`` is imported as-is from ``BurgersFD_CleanFine`` with a stub
``matplotlib`` (not installed here; the module touches ``plt.rcParams`` at import time,
hypernet2D.py:11,24-28).  Nothing is copied from the reference: only its *outputs* on
small seeded inputs are stored, together with this script.

    python oracle/gen_golden.py            # small cases (about a minute)
    python oracle/gen_golden.py --large    # adds the first steps of 250^2 and 750^2

Fixtures (all float64 / int64, ``np.savez_compressed``):
  ops_n12.npz        residual + dense Jacobians (exact and the LSPG rows-only operator)
  fom_n{16,24,48}.npz FOM trajectories + Newton iteration counts
  lspg_n24.npz       POD basis, LSPG trajectory, per-step GN iteration counts
  ecsw_n24.npz       NNLS weights, sample/augmented index sets, HPROM reduced coordinates
  ecswC_n12.npz      ECSW training matrix
  ecm_n24.npz        EmpiricalCubatureMethod selection (indices, weights) on a 24^2 training matrix
  manifold_n24.npz   neural-manifold HPROM (inviscid_burgers_rnm2D_ecsw) with a seeded MLP decoder
  fom_n250.npz, fom_n750.npz   (--large) per-snapshot norms + strided samples
"""
import argparse
import contextlib
import io
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "..", "tests", "golden")
REF_DIR = "/root/reference/BurgersFD_CleanFine"


def import_reference():
    class _RC(dict):
        def update(self, *a, **k):
            pass
    mpl = types.ModuleType("matplotlib")
    plt = types.ModuleType("matplotlib.pyplot")
    plt.rcParams = _RC()
    plt.rc = lambda *a, **k: None
    mpl.pyplot = plt
    sys.modules.setdefault("matplotlib", mpl)
    sys.modules.setdefault("matplotlib.pyplot", plt)
    sys.path.insert(0, REF_DIR)
    import hypernet2D  # noqa: E402  (the reference, unmodified)
    return hypernet2D


@contextlib.contextmanager
def quiet():
    with contextlib.redirect_stdout(io.StringIO()):
        yield


def grid_and_w0(ref, n):
    gx, gy = ref.make_2D_grid(0, 100, 0, 100, n, n)
    return gx, gy, np.ones(2 * n * n)


def record_counts(ref, name):
    """Wrap a reference solver so the per-call iteration counts are recorded."""
    counts = []
    orig = getattr(ref, name)

    def wrapped(*a, **k):
        out = orig(*a, **k)
        counts.append(len(out[1]))
        return out
    setattr(ref, name, wrapped)
    return counts, lambda: setattr(ref, name, orig)


def gen_ops(ref):
    n = 12
    rng = np.random.default_rng(20261104)
    gx, gy, _ = grid_and_w0(ref, n)
    w = 1.0 + rng.random(2 * n * n) * 3.0
    wp = 1.0 + rng.random(2 * n * n) * 3.0
    mu, dt = [4.9, 0.021], 0.05
    Dx, Dy, JDx, JDy, Eye = ref.get_ops(gx, gy)
    r_alt = ref.inviscid_burgers_res2D_alt(w, gx, gy, dt, wp, mu, JDx, JDy)
    r_std = ref.inviscid_burgers_res2D(w, gx, gy, dt, wp, mu, Dx, Dy)
    J = ref.inviscid_burgers_exact_jac2D(w, dt, JDx, JDy, Eye).toarray()
    # operator of inviscid_burgers_implicit2D_LSPG, hypernet2D.py:164-168
    import scipy.sparse as sp
    JDy_q = sp.kron(sp.eye(n), Dy).tocsr()[np.arange(n ** 2).reshape(n, -1).T.flatten(), :]
    Jq = ref.inviscid_burgers_exact_jac2D(w, dt, JDx, JDy_q, Eye).toarray()
    np.savez_compressed(os.path.join(OUT, "ops_n12.npz"), n=n, w=w, wp=wp, mu=mu, dt=dt,
                        r_alt=r_alt, r_std=r_std, J_exact=J, J_lspg=Jq)


def gen_fom(ref, n, steps, mu):
    gx, gy, w0 = grid_and_w0(ref, n)
    counts, restore = record_counts(ref, "newton_raphson")
    with quiet():
        snaps = ref.inviscid_burgers_implicit2D(gx, gy, w0, 0.05, steps, mu)
    restore()
    np.savez_compressed(os.path.join(OUT, f"fom_n{n}.npz"), n=n, mu=mu, dt=0.05, steps=steps,
                        snaps=snaps, newton_its=np.array(counts))
    return snaps


def gen_lspg_ecsw(ref):
    n, steps, k, dt = 24, 12, 6, 0.05
    gx, gy, w0 = grid_and_w0(ref, n)
    train = [[4.25, 0.015], [5.5, 0.03], [4.25, 0.0225]]
    with quiet():
        S = [ref.inviscid_burgers_implicit2D(gx, gy, w0, dt, steps, m) for m in train]
    basis, sigma = ref.POD(np.hstack(S), method="svd")
    basis = np.ascontiguousarray(basis[:, :k])
    mu = [4.75, 0.02]

    counts, restore = record_counts(ref, "gauss_newton_LSPG")
    with quiet():
        snaps, times = ref.inviscid_burgers_implicit2D_LSPG(gx, gy, w0, dt, steps, mu, basis)
    restore()
    np.savez_compressed(os.path.join(OUT, "lspg_n24.npz"), n=n, mu=mu, dt=dt, steps=steps,
                        basis=basis, snaps=snaps, gn_its=np.array(counts), num_its=times[0])

    # ECSW weights the way run_HPROM_ecsw_joshua.py:61-99 builds them (single NNLS block)
    from scipy.optimize import nnls
    tr = S[2]
    with quiet():
        C = ref.compute_ECSW_training_matrix_2D(tr[:, 3:steps:2], tr[:, 0:steps - 3:2], basis,
                                                ref.inviscid_burgers_res2D,
                                                ref.inviscid_burgers_exact_jac2D, gx, gy, dt, train[2])
    idxs = np.zeros((n, n)); idxs[1:-1, 1:-1] = 1
    Ci = np.ascontiguousarray(C[:, (idxs == 1).ravel()])
    wts, _ = nnls(Ci, Ci.sum(axis=1), maxiter=10 ** 9)
    full = 10.0 * np.ones((n, n))
    full[idxs > 0] = wts
    weights = full.ravel()

    sample_inds = np.where(weights != 0)[0]
    aug = ref.generate_augmented_mesh(gx, gy, sample_inds)
    counts, restore = record_counts(ref, "gauss_newton_ECSW_2D")
    with quiet():
        red, _ = ref.inviscid_burgers_ecsw_fixed(gx, gy, weights, w0, dt, steps, mu, basis)
    restore()
    np.savez_compressed(os.path.join(OUT, "ecsw_n24.npz"), n=n, mu=mu, dt=dt, steps=steps,
                        basis=basis, weights=weights, sample_inds=sample_inds.astype(np.int64),
                        augmented_sample=np.asarray(aug, dtype=np.int64), red_coords=red,
                        gn_its=np.array(counts))

    # reduced-mesh operators on a random state (both input modes of res2D_ecsw)
    rng = np.random.default_rng(7)
    wfull = 1.0 + rng.random(2 * n * n)
    wpfull = 1.0 + rng.random(2 * n * n)
    Dx, Dy, JDx, JDy, Eye = ref.get_ops(gx, gy)
    JDx_s = JDx.tolil()[sample_inds, :][:, aug].tocsr()
    JDy_s = JDy.tolil()[sample_inds, :][:, aug].tocsr()
    import scipy.sparse as sp
    E = sp.identity(n * n).tocsr()[sample_inds, :][:, aug]
    E2 = sp.bmat([[E, None], [None, E]]).tocsr()
    idx = np.concatenate((aug, n * n + aug))
    r_full = ref.inviscid_burgers_res2D_ecsw(wfull, gx, gy, dt, wpfull, mu, JDx_s, JDy_s, sample_inds, aug)
    r_red = ref.inviscid_burgers_res2D_ecsw(wfull[idx], gx, gy, dt, wpfull[idx], mu, JDx_s, JDy_s, sample_inds, aug)
    J_red = ref.inviscid_burgers_exact_jac2D_ecsw(wfull[idx], dt, JDx_s, JDy_s, E2, sample_inds, aug).toarray()
    np.savez_compressed(os.path.join(OUT, "ecsw_ops_n24.npz"), n=n, mu=mu, dt=dt, w=wfull, wp=wpfull,
                        weights=weights, r_full=r_full, r_red=r_red, J_red=J_red)


def gen_training_matrix(ref):
    n, steps, k, dt = 12, 8, 4, 0.05
    gx, gy, w0 = grid_and_w0(ref, n)
    mu = [4.25, 0.0225]
    with quiet():
        S = ref.inviscid_burgers_implicit2D(gx, gy, w0, dt, steps, mu)
    basis = np.ascontiguousarray(np.linalg.svd(S, full_matrices=False)[0][:, :k])
    with quiet():
        C = ref.compute_ECSW_training_matrix_2D(S[:, 3:steps:2], S[:, 0:steps - 3:2], basis,
                                                ref.inviscid_burgers_res2D,
                                                ref.inviscid_burgers_exact_jac2D, gx, gy, dt, mu)
    np.savez_compressed(os.path.join(OUT, "ecswC_n12.npz"), n=n, mu=mu, dt=dt, steps=steps,
                        snaps=S, basis=basis, C=C)


def manifold_decoder(q=6, nbar=4, seed=0):
    """The tiny seeded decoder network of the manifold-HPROM fixture (float32, as the reference's)."""
    import torch
    torch.manual_seed(seed)
    net = torch.nn.Sequential(torch.nn.Linear(q + 2, 10), torch.nn.ELU(), torch.nn.Linear(10, nbar))
    with torch.no_grad():
        net[2].weight.mul_(0.05)
        net[2].bias.mul_(0.05)
    return net


def gen_manifold(ref):
    """inviscid_burgers_rnm2D_ecsw (hypernet2D.py:556-674) on the 24^2 ECSW fixture's weights and basis,
    with four further POD modes as the second basis and a seeded MLP as the decoder network."""
    import torch
    g = np.load(os.path.join(OUT, "ecsw_n24.npz"))
    n, steps, dt, mu = int(g["n"]), 8, float(g["dt"]), [4.75, 0.02]
    gx, gy, w0 = grid_and_w0(ref, n)
    train = [[4.25, 0.015], [5.5, 0.03], [4.25, 0.0225]]
    with quiet():
        S = [ref.inviscid_burgers_implicit2D(gx, gy, w0, dt, 12, m) for m in train]
    u, _ = ref.POD(np.hstack(S), method="svd")
    q = g["basis"].shape[1]
    assert np.allclose(np.abs(u[:, :q].T @ g["basis"]), np.eye(q), atol=1e-8)
    basis2 = np.ascontiguousarray(u[:, q:q + 4])
    net = manifold_decoder(q, 4)
    counts, restore = record_counts(ref, "gauss_newton_rnm_ecsw")
    with quiet():
        red, times = ref.inviscid_burgers_rnm2D_ecsw(gx, gy, w0, dt, steps, mu, net, None,
                                                     torch.tensor(g["basis"], dtype=torch.float),
                                                     torch.tensor(basis2, dtype=torch.float), g["weights"])
    restore()
    np.savez_compressed(os.path.join(OUT, "manifold_n24.npz"), n=n, mu=mu, dt=dt, steps=steps, basis=g["basis"],
                        basis2=basis2, weights=g["weights"], red_coords=red, gn_its=np.array(counts), num_its=times[0],
                        **{"net_" + k.replace(".", "_"): v.numpy() for k, v in net.state_dict().items()})


def gen_ecm(ref):
    """EmpiricalCubatureMethod (empirical_cubature_method.py:15-294) driven as run_HPROM_ecm.py:55-91 does, on
    the 24^2 fixture: training matrix of the ECSW training parameter, interior cells, basis of C^T."""
    from empirical_cubature_method import EmpiricalCubatureMethod          # the reference's, unmodified
    g = np.load(os.path.join(OUT, "ecsw_n24.npz"))
    n, steps, dt, mu = int(g["n"]), 12, float(g["dt"]), [4.25, 0.0225]
    gx, gy, w0 = grid_and_w0(ref, n)
    with quiet():
        S = ref.inviscid_burgers_implicit2D(gx, gy, w0, dt, steps, mu)
        C = ref.compute_ECSW_training_matrix_2D(S[:, 3:steps:2], S[:, 0:steps - 3:2], g["basis"],
                                                ref.inviscid_burgers_res2D, ref.inviscid_burgers_exact_jac2D, gx, gy, dt, mu)
    idxs = np.zeros((n, n)); idxs[1:-1, 1:-1] = 1
    C = np.ascontiguousarray(C[:, (idxs == 1).ravel()])
    u, sv, _ = np.linalg.svd(C.T, full_matrices=False)
    u = np.ascontiguousarray(u[:, sv > 1e-8 * sv[0]])
    sel = EmpiricalCubatureMethod()
    with quiet():
        sel.SetUp(u, InitialCandidatesSet=None, constrain_sum_of_weights=True, constrain_conditions=False)
        sel.Run()
    weights = np.zeros(C.shape[1])
    weights[sel.z] = sel.w
    np.savez_compressed(os.path.join(OUT, "ecm_n24.npz"), n=n, C=C, u=u, z=np.asarray(sel.z, dtype=np.int64),
                        w=np.squeeze(sel.w), weights=weights, nerror=sel.nerror,
                        residual=np.linalg.norm(C @ weights - C.sum(axis=1)) / np.linalg.norm(C.sum(axis=1)))


def gen_large(ref, n, steps, mu, stride):
    gx, gy, w0 = grid_and_w0(ref, n)
    counts, restore = record_counts(ref, "newton_raphson")
    with quiet():
        snaps = ref.inviscid_burgers_implicit2D(gx, gy, w0, 0.05, steps, mu)
    restore()
    np.savez_compressed(os.path.join(OUT, f"fom_n{n}.npz"), n=n, mu=mu, dt=0.05, steps=steps,
                        stride=stride, norms=np.linalg.norm(snaps, axis=0),
                        sample=snaps[::stride, :].copy(), newton_its=np.array(counts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--large", action="store_true")
    args = ap.parse_args()
    os.makedirs(OUT, exist_ok=True)
    ref = import_reference()
    gen_ops(ref)
    gen_fom(ref, 16, 20, [4.75, 0.02])
    gen_fom(ref, 24, 15, [5.19, 0.026])
    gen_fom(ref, 48, 10, [4.25, 0.015])
    gen_lspg_ecsw(ref)
    gen_training_matrix(ref)
    gen_manifold(ref)
    gen_ecm(ref)
    if args.large:
        gen_large(ref, 250, 3, [5.19, 0.026], 7)
        gen_large(ref, 750, 2, [4.75, 0.02], 61)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
