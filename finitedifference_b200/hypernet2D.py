"""Drop-in for the hot path of the reference's ``hypernet2D.py``.

Same function names, positional signatures, array layouts and return tuples as
SADPR/FiniteDifference ``BurgersFD_Clean{Coarse,Fine,TestAE}/hypernet2D.py`` for the implicit
2D inviscid-Burgers steppers, so ``run_fom.py`` / ``run_prom.py`` / ``run_HPROM*.py`` can do
``from finitedifference_b200.hypernet2D import ...`` unchanged.  Inputs are NumPy arrays and are
never mutated; outputs are fresh NumPy arrays.  The work runs in hand-written sm_100a kernels
behind the C ABI of include/burgers_b200.h; without the built extension or without a CUDA
device every solver here raises -- there is no CPU path.

Extra, keyword-only arguments (device, strict, ...) default to the reference's
behaviour.
"""
from __future__ import annotations

import glob
import os
import time

import numpy as np

from .discretisation import make_2D_grid  # noqa: F401  (re-exported, hypernet2D.py:2501)
from .fom import FomSolver
from .operators import (get_ops, inviscid_burgers_exact_jac2D, inviscid_burgers_exact_jac2D_ecsw,  # noqa: F401
                        inviscid_burgers_res2D, inviscid_burgers_res2D_alt, inviscid_burgers_res2D_ecsw, make_ddx)
from .ecm import EmpiricalCubatureMethod, ecm_weights  # noqa: F401  (empirical_cubature_method.py:15, run_HPROM_ecm.py:83)
from .nnls import ecsw_weights, nnls  # noqa: F401  (scipy.optimize.nnls at run_HPROM_ecsw_joshua.py:83)
from .pod import POD  # noqa: F401  (hypernet2D.py:2746)
from .rom import (EcswSolver, LspgSolver, ManifoldEcswSolver, ecsw_batch, ecsw_training_matrix,  # noqa: F401
                  generate_augmented_mesh)

__all__ = [
    "make_2D_grid", "inviscid_burgers_implicit2D", "inviscid_burgers_implicit2D_batch",
    "inviscid_burgers_implicit2D_LSPG", "inviscid_burgers_ecsw_fixed", "inviscid_burgers_ecsw_fixed_batch", "generate_augmented_mesh",
    "inviscid_burgers_rnm2D_ecsw", "inviscid_burgers_manifold_ecsw", "ManifoldEcswSolver",
    "compute_ECSW_training_matrix_2D",
    "make_ddx", "get_ops", "inviscid_burgers_res2D", "inviscid_burgers_res2D_alt", "inviscid_burgers_res2D_ecsw",
    "inviscid_burgers_exact_jac2D", "inviscid_burgers_exact_jac2D_ecsw",
    "POD", "nnls", "ecsw_weights", "EmpiricalCubatureMethod", "ecm_weights", "load_or_compute_snaps", "compute_snaps_to_cache", "param_to_snap_fn", "get_saved_params", "compute_error",
]


def inviscid_burgers_implicit2D(grid_x, grid_y, w0, dt, num_steps, mu, *, device=None,
                                strict=True, return_info=False, **solver_kw):
    """Full-order trajectory; reference hypernet2D.py:72-131.

    Returns ``snaps`` of shape ``(2N, num_steps+1)``, float64, C order: column ``i`` is the
    state after ``i`` trapezoidal steps, Newton-converged to a relative residual of 1e-12
    (at most 100 iterations) exactly as ``newton_raphson`` decides it.  ``return_info=True``
    also returns ``{"newton_its": ..., "relnorm": ...}`` per step.
    """
    solver = FomSolver(grid_x, grid_y, dt, num_steps, [list(mu)[:2]], device=device, strict=strict,
                       **solver_kw)
    solver.set_initial(np.asarray(w0, dtype=np.float64).ravel())
    solver.run()
    snaps = solver.snapshots(0)
    if return_info:
        its, rel = solver.diagnostics()
        return snaps, {"newton_its": its[0], "relnorm": rel[0], "launches": solver.launches,
                       "steps_redone": solver.steps_redone}
    return snaps


def inviscid_burgers_implicit2D_batch(grid_x, grid_y, w0, dt, num_steps, mus, *, device=None,
                                      strict=True, **solver_kw):
    """Parameter sweep: the FOM for every ``mu`` in ``mus`` advanced together on one GPU.

    Not in the reference (which loops ``load_or_compute_snaps`` over mu serially,
    run_prom.py:59-71); returns the ready-to-export ``FomSolver`` so that callers choose
    between ``snapshots(m)`` (host) and ``snapshots_device(m)`` (HBM-resident).
    """
    solver = FomSolver(grid_x, grid_y, dt, num_steps, mus, device=device, strict=strict, **solver_kw)
    solver.set_initial(w0)
    solver.run()
    return solver


def inviscid_burgers_implicit2D_LSPG(grid_x, grid_y, w0, dt, num_steps, mu, basis, *, device=None,
                                     jacobian="reference", return_info=False):
    """LSPG PROM trajectory; reference hypernet2D.py:133-200.

    Returns ``(snaps (2N, T+1), (num_its, jac_time, res_time, ls_time))`` like the reference.
    ``jacobian="reference"`` (default) reproduces the operator the reference builds at :167;
    ``"exact"`` uses the true Jacobian.  ``return_info=True`` appends
    ``{"red_coords": (k, T+1), "gn_its": per-step residual evaluations}``.
    """
    solver = LspgSolver(grid_x, grid_y, dt, mu, np.asarray(basis, dtype=np.float64), device=device,
                        jacobian=jacobian)
    snaps, red, hist = solver.run(np.asarray(w0, dtype=np.float64).ravel(), int(num_steps))
    out = (np.ascontiguousarray(snaps.T.cpu().numpy()),
           (sum(len(h) for h in hist), solver.jac_time, solver.res_time, solver.ls_time))
    if return_info:
        out += ({"red_coords": np.ascontiguousarray(red.T.cpu().numpy()),
                 "gn_its": np.array([len(h) for h in hist]), "resnorms": hist},)
    return out


def inviscid_burgers_ecsw_fixed(grid_x, grid_y, weights, w0, dt, num_steps, mu, basis, *, device=None,
                                return_info=False):
    """ECSW/ECM HPROM trajectory on the sampled mesh; reference hypernet2D.py:202-273.

    Returns ``(red_coords (k, T+1), (jac_time, res_time, ls_time))`` like the reference.
    ``return_info=True`` appends ``{"sample_inds", "augmented_sample", "gn_its"}``.
    """
    solver = EcswSolver(grid_x, grid_y, weights, dt, mu, np.asarray(basis, dtype=np.float64), device=device)
    red, hist = solver.run(np.asarray(w0, dtype=np.float64).ravel(), int(num_steps))
    out = (np.ascontiguousarray(red.T.cpu().numpy()), (solver.jac_time, solver.res_time, solver.ls_time))
    if return_info:
        out += ({"sample_inds": solver.sample_inds, "augmented_sample": solver.augmented_sample,
                 "gn_its": np.array([len(h) for h in hist]), "resnorms": hist},)
    return out


def inviscid_burgers_ecsw_fixed_batch(grid_x, grid_y, weights, w0, dt, num_steps, mus, basis, *, device=None):
    """``inviscid_burgers_ecsw_fixed`` for a list of test parameters (the loops over test mu of
    run_tests.py:9-10 / run_HPROM_ecsw_joshua.py): the HPROMs run concurrently on one GPU.
    Returns a list of ``red_coords (k, T+1)`` arrays in the order of ``mus``."""
    return [red for red, _ in ecsw_batch(grid_x, grid_y, weights, w0, dt, num_steps, mus, basis, device=device)]


def inviscid_burgers_manifold_ecsw(grid_x, grid_y, weights, y0, dt, num_steps, mu, decode, decode_jac, *, device=None,
                                   max_its=20, relnorm_cutoff=1e-5, min_delta=0.1, return_info=False):
    """Sampled-mesh HPROM with a caller-supplied decoder: the common core of the reference's manifold
    ROMs (``inviscid_burgers_rnm2D_ecsw`` :556-674, ``_pod_rbf_2D_global_ecsw`` :973-1111,
    ``_pod_gp_2D_ecsw`` :1113-1275).  ``decode(y) -> (2 n_aug,)`` and ``decode_jac(y) -> (2 n_aug, q)`` are
    callables on CUDA tensors (entries ``ManifoldEcswSolver.idx`` of the full state); ``max_its=10``
    reproduces ``gauss_newton_pod_rbf_ecsw`` (:2293).  Returns ``(red_coords (q, T+1), (num_its, jac_time,
    res_time, ls_time))``."""
    import torch
    solver = ManifoldEcswSolver(grid_x, grid_y, weights, dt, mu, device=device)
    y0 = y0 if isinstance(y0, torch.Tensor) else torch.from_numpy(np.asarray(y0, dtype=np.float64))
    red, hist = solver.run(y0.to(solver.device), int(num_steps), decode, decode_jac, max_its=max_its,
                           relnorm_cutoff=relnorm_cutoff, min_delta=min_delta)
    out = (np.ascontiguousarray(red.T.double().cpu().numpy()),
           (sum(len(h) for h in hist), solver.jac_time, solver.res_time, solver.ls_time))
    if return_info:
        out += ({"gn_its": np.array([len(h) for h in hist]), "resnorms": hist, "idx": solver.idx,
                 "sample_inds": solver.sample_inds, "augmented_sample": solver.augmented_sample},)
    return out


def inviscid_burgers_rnm2D_ecsw(grid_x, grid_y, w0, dt, num_steps, mu, rnm, ref, basis, basis2, weights, *,
                                device=None, return_info=False):
    """Neural-manifold HPROM, reference hypernet2D.py:556-674, same positional signature: the state is
    ``basis @ y + basis2 @ rnm(cat(y, mu))``.  As in the reference the decoder and ``y`` are float32
    (``dtype=torch.float``, :592-596, :2214) and the decoder Jacobian comes from forward-mode autodiff
    (``functorch.jacfwd``, :613); residual, ``J @ dW/dy``, normal equations and the solve are float64
    kernels.  ``ref`` is unused there and here.  Returns ``(red_coords (q, T+1), (num_its, jac_time,
    res_time, ls_time))``."""
    import torch
    solver = ManifoldEcswSolver(grid_x, grid_y, weights, dt, mu, device=device)
    dev = solver.device
    f32 = dict(dtype=torch.float32, device=dev)
    tmu = torch.tensor(list(mu)[:2], **f32)
    B1 = torch.as_tensor(basis).to(**f32)
    B2 = torch.as_tensor(basis2).to(**f32)
    net = rnm.to(dev)
    w0t = torch.as_tensor(np.asarray(w0, dtype=np.float64).ravel()).to(**f32)
    y0 = B1.T @ w0t                                                      # :593
    idx = torch.from_numpy(solver.idx).to(dev)
    V, Vbar = B1[idx].contiguous(), B2[idx].contiguous()                 # :601-602

    def decode32(x):
        return V @ x + Vbar @ net(torch.cat((x, tmu)))                   # :603-605

    def decode(x):
        with torch.no_grad():
            return decode32(x)

    jac = torch.func.jacfwd(decode32)                                    # :613

    def decode_jac(x):
        return jac(x).detach()

    def update(y, dy):
        return y + dy.to(torch.float32)                                  # :2214

    red, hist = solver.run(y0, int(num_steps), decode, decode_jac, update=update)
    out = (np.ascontiguousarray(red.T.double().cpu().numpy()),
           (sum(len(h) for h in hist), solver.jac_time, solver.res_time, solver.ls_time))
    if return_info:
        out += ({"gn_its": np.array([len(h) for h in hist]), "resnorms": hist},)
    return out


_RES_NAMES = ("inviscid_burgers_res2D", "inviscid_burgers_res2D_alt")
_JAC_NAMES = ("inviscid_burgers_exact_jac2D",)


def _check_callable(fn, names, what):
    """The device kernels evaluate the residual / exact Jacobian of ``get_ops`` themselves, so the only
    callables that can be honoured are this module's or the reference's functions of those names."""
    if fn is None:
        return
    name = getattr(fn, "__name__", None)
    if name not in names:
        raise ValueError(f"compute_ECSW_training_matrix_2D: {what}={name or fn!r} is not supported; the device path "
                         f"evaluates {' / '.join(names)} (pass that function or None)")


def compute_ECSW_training_matrix_2D(snaps, prev_snaps, basis, res=None, jac=None, grid_x=None, grid_y=None,
                                    dt=None, mu=None, *, device=None):
    """ECSW training matrix ``C`` (k*n_snaps, N); reference hypernet2D.py:2795-2816.

    Same positional signature as the reference.  ``res`` and ``jac`` must be ``inviscid_burgers_res2D``
    (or ``_res2D_alt``: same values) and ``inviscid_burgers_exact_jac2D`` -- this module's or the
    reference's (run_HPROM_ecsw_joshua.py:61-64 passes exactly these) -- or None: the residual and the
    exact Jacobian of ``get_ops`` are evaluated by the device kernels.  Any other callable raises
    ``ValueError`` instead of silently producing the matrix of a different operator.
    """
    _check_callable(res, _RES_NAMES, "res")
    _check_callable(jac, _JAC_NAMES, "jac")
    return ecsw_training_matrix(snaps, prev_snaps, basis, grid_x, grid_y, dt, mu, device=device)


# -- snapshot cache (hypernet2D.py:3150-3221) ------------------------------------------------
# param_to_snap_fn, get_saved_params, load_or_compute_snaps and compute_error below are RESTATED from
# the reference (hypernet2D.py:3150-3221) statement by statement, print strings included: the cache's
# file names, the messages scripts grep for and the error definition are the observable behaviour a
# drop-in must keep, so there is nothing to redesign in these ~35 lines.
def param_to_snap_fn(mu, snap_folder="param_snaps", suffix=".npy"):
    """``param_snaps/mu1_<str(mu1)>+mu2_<str(mu2)>.npy`` -- same names as the reference."""
    npar = len(mu)
    snapfn = snap_folder + "/"
    for i in range(npar):
        if i > 0:
            snapfn += "+"
        snapfn += "mu{}_{}".format(i + 1, mu[i])
    return snapfn + suffix


def get_saved_params(snap_folder="param_snaps"):
    return set(glob.glob(snap_folder + "/*"))


def load_or_compute_snaps(mu, grid_x, grid_y, w0, dt, num_steps, snap_folder="param_snaps"):
    """Load ``(2N, T+1)`` snapshots from the mu-keyed ``.npy`` cache or compute and save them.

    File names and contents are interchangeable with the reference's cache.
    """
    if not os.path.exists(snap_folder):
        os.makedirs(snap_folder)
    snap_fn = param_to_snap_fn(mu, snap_folder=snap_folder)
    if snap_fn in get_saved_params(snap_folder=snap_folder):
        print(f"Loading saved snaps for mu1={mu[0]}, mu2={mu[1]}")
        snaps = np.load(snap_fn)[:, :num_steps + 1]
    else:
        print(f"Computing new snaps for mu1={mu[0]}, mu2={mu[1]}")
        t0 = time.time()
        snaps = inviscid_burgers_implicit2D(grid_x, grid_y, w0, dt, num_steps, mu)
        print("Elapsed time: {:3.3e}".format(time.time() - t0))
        np.save(snap_fn, snaps)
    return snaps


def compute_snaps_to_cache(mus, grid_x, grid_y, w0, dt, num_steps, snap_folder="param_snaps", *, device=None,
                           **sweep_kw):
    """``load_or_compute_snaps`` for a list of parameters: the missing trajectories are computed
    together (``SnapshotSweep``: groups of samples on worker streams) and written by the worker
    threads as they reach the host, under the reference's file names.  Returns the file names, in
    the order of ``mus``; ``np.load(fn, mmap_mode="r")`` gives the ``(2N, T+1)`` array."""
    from .sweep import SnapshotSweep, save_sink
    if not os.path.exists(snap_folder):
        os.makedirs(snap_folder)
    names = [param_to_snap_fn(list(mu), snap_folder=snap_folder) for mu in mus]
    saved = get_saved_params(snap_folder=snap_folder)
    todo = [i for i, fn in enumerate(names) if fn not in saved]
    if todo:
        sweep = SnapshotSweep(grid_x, grid_y, dt, num_steps, device=device, **sweep_kw)
        sweep.run(w0, [mus[i] for i in todo], sink=save_sink(snap_folder))
    return names


def compute_error(rom_snaps, hdm_snaps):
    """Relative error at each time step and its mean (hypernet2D.py:3150-3155)."""
    sq_hdm = np.sqrt(np.square(rom_snaps).sum(axis=0))
    sq_err = np.sqrt(np.square(rom_snaps - hdm_snaps).sum(axis=0))
    rel_err = sq_err / sq_hdm
    return rel_err, rel_err.mean()
