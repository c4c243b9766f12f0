"""CPU oracle for the implicit 2D parametrised inviscid-Burgers hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package may import this module:
only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl
reference`` legs of ``bench.py`` are allowed to, and there only as the checker or as
the timed CPU baseline -- never as the thing shipped.

It restates, in NumPy/SciPy, the algorithm of SADPR/FiniteDifference's
``BurgersFD_CleanFine/hypernet2D.py`` (the hot-path functions are byte-identical in
the Coarse / TestAE copies).  Each function cites the reference lines it follows.
The arithmetic is the reference's: the residual is evaluated term by term in the same
order, the Newton correction is obtained from the assembled sparse Jacobian with
SuperLU (``scipy.sparse.linalg.spsolve``) and the reduced problems with LAPACK
``gelsd`` (``numpy.linalg.lstsq``), exactly the third-party calls the reference makes.
The sparse operators are assembled directly from the stencil instead of through
``kron``/``bmat``/fancy indexing, which is the only structural difference.

Parity pin: the reference ships no golden vectors or tests for this path
(SURVEY.md section 4), so this oracle is pinned against outputs of the reference
itself, run unmodified in the build container by ``oracle/gen_golden.py`` and
committed under ``tests/golden/`` (``tests/test_oracle_golden.py`` replays them).
"""
from __future__ import annotations

import time

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

__all__ = [
    "make_2D_grid", "backward_difference", "lifted_operators", "residual",
    "jacobian", "newton_raphson", "fom_trajectory", "gauss_newton_lspg",
    "lspg_trajectory", "augmented_mesh", "ecsw_operators", "residual_ecsw",
    "jacobian_ecsw", "gauss_newton_ecsw", "ecsw_trajectory",
    "ecsw_training_matrix", "compute_error", "param_to_snap_fn",
]


# --------------------------------------------------------------------------------------
# grids and difference operators
# --------------------------------------------------------------------------------------
def make_2D_grid(x_low, x_up, y_low, y_up, num_cells_x, num_cells_y):
    """Cell-edge coordinates (hypernet2D.py:2501-2507)."""
    return (np.linspace(x_low, x_up, num_cells_x + 1),
            np.linspace(y_low, y_up, num_cells_y + 1))


def backward_difference(grid):
    """First-order upwind difference matrix of ``make_ddx`` (hypernet2D.py:2486-2492).

    ``A[i, i] = 1/dx[i]`` and ``A[i, i-1] = -1/dx[i-1]``: the sub-diagonal entry is
    scaled by the spacing of the *upwind* cell, as ``spdiags`` places ``data[k][j]`` in
    column ``j``.  Row 0 has only its diagonal -- a missing neighbour contributes 0.
    """
    d = grid[1:] - grid[:-1]
    n = d.size
    rows = np.concatenate((np.arange(n), np.arange(1, n)))
    cols = np.concatenate((np.arange(n), np.arange(0, n - 1)))
    vals = np.concatenate((1.0 / d, -1.0 / d[:-1]))
    return sp.csr_matrix((vals, (rows, cols)), shape=(n, n))


def lifted_operators(grid_x, grid_y, y_columns="permuted"):
    """N x N operators acting on a row-major (y, x) field: ``get_ops`` (hypernet2D.py:2509-2520).

    ``JDx`` differences along x (``kron(I_ny, Dx)``, :2512).  ``JDy`` is
    ``kron(I_nx, Dy)`` with rows and columns permuted from column-major to row-major
    (:2513-2518), i.e. a difference along y.

    ``y_columns="unpermuted"`` reproduces the operator that
    ``inviscid_burgers_implicit2D_LSPG`` builds at hypernet2D.py:167, where only the
    rows are permuted: row (r, c) then reads cells (c, r) and (c, r-1).  The LSPG PROM
    Jacobian of the reference is built from that operator.
    """
    nx, ny = grid_x.size - 1, grid_y.size - 1
    n_cells = nx * ny
    dx = grid_x[1:] - grid_x[:-1]
    dy = grid_y[1:] - grid_y[:-1]
    r, c = np.divmod(np.arange(n_cells), nx)

    has_l = c >= 1
    rows = np.concatenate((np.arange(n_cells), np.nonzero(has_l)[0]))
    cols = np.concatenate((np.arange(n_cells), np.nonzero(has_l)[0] - 1))
    vals = np.concatenate((1.0 / dx[c], -1.0 / dx[c[has_l] - 1]))
    jdx = sp.csr_matrix((vals, (rows, cols)), shape=(n_cells, n_cells))

    has_d = r >= 1
    rows = np.concatenate((np.arange(n_cells), np.nonzero(has_d)[0]))
    vals = np.concatenate((1.0 / dy[r], -1.0 / dy[r[has_d] - 1]))
    if y_columns == "permuted":
        cols = np.concatenate((np.arange(n_cells), np.nonzero(has_d)[0] - nx))
    elif y_columns == "unpermuted":
        if nx != ny:
            raise ValueError("the reference's LSPG operator (hypernet2D.py:167) needs nx == ny")
        # column index of kron(I_nx, Dy) is c*ny + r, read as a row-major cell id
        cols = np.concatenate((c * ny + r, c[has_d] * ny + r[has_d] - 1))
    else:
        raise ValueError(y_columns)
    jdy = sp.csr_matrix((vals, (rows, cols)), shape=(n_cells, n_cells))
    jdx.sort_indices()
    jdy.sort_indices()
    return jdx, jdy


# --------------------------------------------------------------------------------------
# full-order residual and Jacobian
# --------------------------------------------------------------------------------------
def _shift_x(f):
    out = np.zeros_like(f)
    out[:, 1:] = f[:, :-1]
    return out


def _shift_y(f):
    out = np.zeros_like(f)
    out[1:, :] = f[:-1, :]
    return out


def residual(w, grid_x, grid_y, dt, wp, mu):
    """Trapezoidal-rule residual of one time step.

    Follows ``inviscid_burgers_res2D_alt`` (hypernet2D.py:2588-2646) term by term;
    ``inviscid_burgers_res2D`` (:2544-2586) is the same arithmetic on reshaped fields.
    The operator products ``(0.5*dt*JDx) @ F`` are evaluated matrix-free as
    ``cx*F[p] + (-cxl)*F[L]`` with the pre-scaled coefficients the reference's
    ``0.5 * dt * JDxec`` holds.
    """
    nx, ny = grid_x.size - 1, grid_y.size - 1
    dx = grid_x[1:] - grid_x[:-1]
    dy = grid_y[1:] - grid_y[:-1]
    xc = (grid_x[1:] + grid_x[:-1]) / 2
    u, v = (a.reshape(ny, nx) for a in np.split(np.asarray(w, dtype=np.float64), 2))
    up, vp = (a.reshape(ny, nx) for a in np.split(np.asarray(wp, dtype=np.float64), 2))

    fx = 0.5 * np.square(u) + 0.5 * np.square(up)        # :2619
    hy = 0.5 * np.square(v) + 0.5 * np.square(vp)        # :2620
    g = 0.5 * u * v + 0.5 * up * vp                      # :2621

    half_dt = 0.5 * dt
    cx = (half_dt * (1.0 / dx))[None, :]                 # diagonal of 0.5*dt*JDx
    cxl = np.zeros((1, nx)); cxl[0, 1:] = half_dt * (-(1.0 / dx[:-1]))
    cy = (half_dt * (1.0 / dy))[:, None]
    cyd = np.zeros((ny, 1)); cyd[1:, 0] = half_dt * (-(1.0 / dy[:-1]))

    def ddx(f):
        return cxl * _shift_x(f) + cx * f

    def ddy(f):
        return cyd * _shift_y(f) + cy * f

    src = dt * 0.02 * np.exp(mu[1] * xc)                 # :2625
    ru = u - up + ddx(fx) + ddy(g) - src[None, :]        # :2637-2638
    # :2628-2631, :2639 -- dx broadcast down the first column (uniform grids: dx[r] == dx[0])
    ru[:, 0] -= 0.5 * dt * mu[0] ** 2 / (dx if nx == ny else dx[0])
    rv = v - vp + ddy(hy) + ddx(g)                       # :2642-2643
    return np.concatenate((ru.ravel(), rv.ravel()))


def jacobian(w, dt, jdx, jdy):
    """Exact Jacobian ``[[ul, ur], [ll, lr]] + I`` (hypernet2D.py:2703-2732)."""
    u, v = np.split(np.asarray(w, dtype=np.float64), 2)
    ud = 0.5 * dt * sp.diags(u)
    vd = 0.5 * dt * sp.diags(v)
    ul = jdx @ ud + 0.5 * jdy @ vd
    ur = 0.5 * jdy @ ud
    ll = 0.5 * jdx @ vd
    lr = jdy @ vd + 0.5 * jdx @ ud
    n2 = 2 * u.size
    return (sp.bmat([[ul, ur], [ll, lr]]) + sp.identity(n2)).tocsr()


# --------------------------------------------------------------------------------------
# FOM: Newton-Raphson + time stepping
# --------------------------------------------------------------------------------------
def newton_raphson(func, jac, x0, max_its=20, relnorm_cutoff=1e-12):
    """Newton with a relative-residual stop (hypernet2D.py:1887-1933).

    The test precedes the solve, so the iterate that passes it is returned un-updated;
    ``init_norm`` is the residual norm at ``x0``.  Returns ``(x, resnorms)``.
    """
    x = x0.copy()
    init_norm = np.linalg.norm(func(x0))
    resnorms = []
    for _ in range(max_its):
        resnorm = np.linalg.norm(func(x))
        resnorms.append(resnorm)
        if resnorm / init_norm < relnorm_cutoff:
            break
        J = jac(x)
        f = func(x)
        x -= spla.spsolve(J, f)
    return x, resnorms


def fom_trajectory(grid_x, grid_y, w0, dt, num_steps, mu, return_info=False):
    """``inviscid_burgers_implicit2D`` (hypernet2D.py:72-131).

    Returns ``snaps`` of shape ``(2N, num_steps+1)`` (C order).  With ``return_info``
    also the per-step list of residual-norm histories (their lengths are the Newton
    iteration counts the parity tests compare).
    """
    w0 = np.asarray(w0, dtype=np.float64).ravel()
    snaps = np.zeros((w0.size, num_steps + 1))
    snaps[:, 0] = w0
    wp = w0.copy()
    jdx, jdy = lifted_operators(grid_x, grid_y)
    history = []
    for i in range(num_steps):
        w, resnorms = newton_raphson(
            lambda x: residual(x, grid_x, grid_y, dt, wp, mu),
            lambda x: jacobian(x, dt, jdx, jdy),
            wp, max_its=100, relnorm_cutoff=1e-12)               # :123
        history.append(resnorms)
        snaps[:, i + 1] = w
        wp = w.copy()
    return (snaps, history) if return_info else snaps


# --------------------------------------------------------------------------------------
# LSPG PROM
# --------------------------------------------------------------------------------------
def gauss_newton_lspg(func, jac, basis, y0, max_its=20, relnorm_cutoff=1e-5, min_delta=0.1):
    """``gauss_newton_LSPG`` (hypernet2D.py:1935-2005). Returns ``(y, resnorms, times)``."""
    jac_time = res_time = ls_time = 0.0
    y = y0.copy()
    w = basis.dot(y0)
    init_norm = np.linalg.norm(func(w))
    resnorms = []
    for _ in range(max_its):
        resnorm = np.linalg.norm(func(w))
        resnorms.append(resnorm)
        if resnorm / init_norm < relnorm_cutoff:
            break
        if len(resnorms) > 1 and abs((resnorms[-2] - resnorms[-1]) / resnorms[-2]) < min_delta:
            break
        t0 = time.time(); J = jac(w); jac_time += time.time() - t0
        t0 = time.time(); f = func(w); res_time += time.time() - t0
        t0 = time.time()
        JV = J.dot(basis)
        dy = np.linalg.lstsq(JV, -f, rcond=None)[0]
        ls_time += time.time() - t0
        y += dy
        w = basis.dot(y)
    return y, resnorms, (jac_time, res_time, ls_time)


def lspg_trajectory(grid_x, grid_y, w0, dt, num_steps, mu, basis, jacobian_mode="reference",
                    return_info=False):
    """``inviscid_burgers_implicit2D_LSPG`` (hypernet2D.py:133-200).

    ``jacobian_mode="reference"`` uses the rows-only-permuted y operator of line 167
    (what the reference runs); ``"exact"`` uses the true Jacobian.
    Returns ``(snaps, (num_its, jac_time, res_time, ls_time))``; with ``return_info``
    additionally ``(red_coords, per-step resnorm lists)``.
    """
    basis = np.asarray(basis, dtype=np.float64)
    npod = basis.shape[1]
    w0 = np.asarray(w0, dtype=np.float64).ravel()
    snaps = np.zeros((w0.size, num_steps + 1))
    red = np.zeros((npod, num_steps + 1))
    y0 = basis.T.dot(w0)
    w0 = basis.dot(y0)
    snaps[:, 0], red[:, 0] = w0, y0
    wp, yp = w0.copy(), y0.copy()
    jdx, jdy = lifted_operators(
        grid_x, grid_y, "unpermuted" if jacobian_mode == "reference" else "permuted")
    num_its, jt, rt, lt = 0, 0.0, 0.0, 0.0
    history = []
    for i in range(num_steps):
        y, resnorms, times = gauss_newton_lspg(
            lambda x: residual(x, grid_x, grid_y, dt, wp, mu),
            lambda x: jacobian(x, dt, jdx, jdy), basis, yp)
        num_its += len(resnorms)
        jt += times[0]; rt += times[1]; lt += times[2]
        history.append(resnorms)
        w = basis.dot(y)
        red[:, i + 1], snaps[:, i + 1] = y, w
        wp, yp = w.copy(), y.copy()
    out = (snaps, (num_its, jt, rt, lt))
    return out + (red, history) if return_info else out


# --------------------------------------------------------------------------------------
# ECSW / ECM hyper-reduced PROM on the sampled mesh
# --------------------------------------------------------------------------------------
def augmented_mesh(grid_x, grid_y, sample_inds):
    """``generate_augmented_mesh`` (hypernet2D.py:2522-2542).

    Sampled cells plus their (r-1, c) and (r, c-1) neighbours, sorted ascending, unique,
    int64.  Integer work: parity on this is bit-exact.
    """
    nx = grid_x.size - 1
    s = np.asarray(sample_inds, dtype=np.int64)
    r, c = np.divmod(s, nx)
    return np.unique(np.concatenate((s, s[r >= 1] - nx, s[c >= 1] - 1))).astype(np.int64)


def ecsw_operators(grid_x, grid_y, sample_inds, augmented_sample):
    """Rows ``sample_inds`` x columns ``augmented_sample`` of JDx, JDy and the identity
    (hypernet2D.py:220-236)."""
    jdx, jdy = lifted_operators(grid_x, grid_y)
    n_cells = jdx.shape[0]
    eye = sp.identity(n_cells, format="csr")[sample_inds, :][:, augmented_sample]
    eye2 = sp.bmat([[eye, None], [None, eye]]).tocsr()
    return (jdx[sample_inds, :][:, augmented_sample].tocsr(),
            jdy[sample_inds, :][:, augmented_sample].tocsr(), eye2)


def residual_ecsw(w, grid_x, grid_y, dt, wp, mu, jdx_s, jdy_s, sample_inds, augmented_sample):
    """``inviscid_burgers_res2D_ecsw`` (hypernet2D.py:2648-2701).

    ``w``/``wp`` are either full states (2N) or states restricted to
    ``[aug | N+aug]`` (2 n_aug).  Output rows are ordered ``[sample (u) | sample (v)]``.
    """
    dx = grid_x[1:] - grid_x[:-1]
    dy = grid_y[1:] - grid_y[:-1]
    xc = (grid_x[1:] + grid_x[:-1]) / 2
    nx = dx.size
    lbc = np.where(sample_inds % nx == 0, 0.5 * dt * mu[0] ** 2 / dx[0], 0.0)   # :2660-2665
    src = np.tile(dt * 0.02 * np.exp(mu[1] * xc), dy.size)[sample_inds]         # :2666-2669
    u, v = np.split(np.asarray(w, dtype=np.float64), 2)
    up, vp = np.split(np.asarray(wp, dtype=np.float64), 2)
    if u.size > augmented_sample.size:                                          # :2673
        ua, va, upa, vpa = (a[augmented_sample] for a in (u, v, up, vp))
        us, vs, ups, vps = (a[sample_inds] for a in (u, v, up, vp))
    else:
        ua, va, upa, vpa = u, v, up, vp
        overlap = np.isin(augmented_sample, sample_inds)
        us, vs, ups, vps = (a[overlap] for a in (u, v, up, vp))
    fx = 0.5 * np.square(ua) + 0.5 * np.square(upa)
    hy = 0.5 * np.square(va) + 0.5 * np.square(vpa)
    g = 0.5 * ua * va + 0.5 * upa * vpa
    ru = us - ups + 0.5 * dt * jdx_s @ fx + 0.5 * dt * jdy_s @ g - src
    ru -= lbc
    rv = vs - vps + 0.5 * dt * jdy_s @ hy + 0.5 * dt * (jdx_s @ g)
    return np.concatenate((ru, rv))


def jacobian_ecsw(w, dt, jdx_s, jdy_s, eye2, sample_inds, augmented_sample):
    """``inviscid_burgers_exact_jac2D_ecsw`` (hypernet2D.py:2734-2744): (2 n_s, 2 n_aug) CSR."""
    u, v = np.split(np.asarray(w, dtype=np.float64), 2)
    if u.size > augmented_sample.size:
        u, v = u[augmented_sample], v[augmented_sample]
    ud, vd = 0.5 * dt * sp.diags(u), 0.5 * dt * sp.diags(v)
    ul = jdx_s @ ud + 0.5 * jdy_s @ vd
    ur = 0.5 * jdy_s @ ud
    ll = 0.5 * jdx_s @ vd
    lr = jdy_s @ vd + 0.5 * jdx_s @ ud
    return (sp.bmat([[ul, ur], [ll, lr]]) + eye2).tocsr()


def gauss_newton_ecsw(func, jac, basis, y0, weight, stepsize=1, max_its=20,
                      relnorm_cutoff=1e-5, min_delta=1e-1):
    """``gauss_newton_ECSW_2D`` (hypernet2D.py:2007-2044). Returns ``(y, resnorms, times)``."""
    y = y0.copy()
    w = basis.dot(y0)
    weights = np.concatenate((weight, weight))
    init_norm = np.linalg.norm(func(w) * weights)
    resnorms = []
    jac_time = res_time = ls_time = 0.0
    for _ in range(max_its):
        resnorm = np.linalg.norm(func(w) * weights)
        resnorms.append(resnorm)
        if resnorm / init_norm < relnorm_cutoff:
            break
        if len(resnorms) > 1 and abs((resnorms[-2] - resnorms[-1]) / resnorms[-2]) < min_delta:
            break
        t0 = time.time(); J = jac(w); jac_time += time.time() - t0
        JVw = sp.diags(weights) @ (J @ basis)
        t0 = time.time(); f = func(w); res_time += time.time() - t0
        t0 = time.time()
        dy = np.linalg.lstsq(JVw, -(f * weights), rcond=None)[0]
        ls_time += time.time() - t0
        y = y + stepsize * dy
        w = basis.dot(y)
    return y, resnorms, (jac_time, res_time, ls_time)


def ecsw_trajectory(grid_x, grid_y, weights, w0, dt, num_steps, mu, basis, return_info=False):
    """``inviscid_burgers_ecsw_fixed`` (hypernet2D.py:202-273).

    Returns ``(red_coords (k, T+1), (jac_time, res_time, ls_time))``; with
    ``return_info`` additionally ``(sample_inds, augmented_sample, per-step resnorms)``.
    """
    basis = np.asarray(basis, dtype=np.float64)
    weights = np.asarray(weights, dtype=np.float64)
    w0 = np.asarray(w0, dtype=np.float64).ravel()
    npod = basis.shape[1]
    n_cells = w0.size // 2
    red = np.zeros((npod, num_steps + 1))
    y0 = basis.T.dot(w0)
    w0 = basis.dot(y0)
    red[:, 0] = y0
    yp = y0.copy()
    sample_inds = np.where(weights != 0)[0]                                     # :225
    aug = augmented_mesh(grid_x, grid_y, sample_inds)
    jdx_s, jdy_s, eye2 = ecsw_operators(grid_x, grid_y, sample_inds, aug)
    sample_weights = np.hstack((weights, weights))[sample_inds]                 # :239-240
    idx = np.concatenate((aug, n_cells + aug))
    wp = w0[idx]
    basis_red = basis[idx, :]
    jt = rt = lt = 0.0
    history = []
    for i in range(num_steps):
        y, resnorms, times = gauss_newton_ecsw(
            lambda x: residual_ecsw(x, grid_x, grid_y, dt, wp, mu, jdx_s, jdy_s, sample_inds, aug),
            lambda x: jacobian_ecsw(x, dt, jdx_s, jdy_s, eye2, sample_inds, aug),
            basis_red, yp, sample_weights)
        jt += times[0]; rt += times[1]; lt += times[2]
        history.append(resnorms)
        red[:, i + 1] = y
        wp = basis_red.dot(y)
        yp = y.copy()
    out = (red, (jt, rt, lt))
    return out + (sample_inds, aug, history) if return_info else out


def gauss_newton_manifold_ecsw(func, jac, y0, decode, decode_jac, weight, max_its=20, relnorm_cutoff=1e-5,
                               min_delta=0.1):
    """``gauss_newton_rnm_ecsw`` (hypernet2D.py:2167-2217): Gauss-Newton on reduced coordinates with a
    nonlinear decoder; ``y`` is a float32 torch tensor as in the reference (:2214)."""
    import torch
    y = y0.detach().clone()
    w = decode(y).numpy().astype(np.float64)
    weights = np.concatenate((weight, weight))
    init_norm = np.linalg.norm(func(w) * weights)
    resnorms = []
    for _ in range(max_its):
        resnorm = np.linalg.norm(func(w) * weights)
        resnorms.append(resnorm)
        if resnorm / init_norm < relnorm_cutoff:
            break
        if len(resnorms) > 1 and abs((resnorms[-2] - resnorms[-1]) / resnorms[-2]) < min_delta:
            break
        J = jac(w)
        V = decode_jac(y).numpy()
        fw = func(w) * weights
        JVw = sp.diags(weights) @ (J @ V)
        dy = np.linalg.lstsq(JVw, -fw, rcond=None)[0]
        y = y + torch.tensor(dy, dtype=torch.float)
        w = decode(y).numpy().astype(np.float64)
    return y, resnorms


def manifold_ecsw_trajectory(grid_x, grid_y, w0, dt, num_steps, mu, net, basis, basis2, weights):
    """``inviscid_burgers_rnm2D_ecsw`` (hypernet2D.py:556-674): neural-manifold HPROM on the sampled mesh,
    state = basis @ y + basis2 @ net(cat(y, mu)), float32 decoder, float64 residual / Jacobian.
    Returns ``(red_coords (q, T+1), per-step resnorm lists)``."""
    import torch
    weights = np.asarray(weights, dtype=np.float64)
    n_cells = weights.size
    sample_inds = np.where(weights != 0)[0]
    aug = augmented_mesh(grid_x, grid_y, sample_inds)
    jdx_s, jdy_s, eye2 = ecsw_operators(grid_x, grid_y, sample_inds, aug)
    sample_weights = np.concatenate((weights, weights))[sample_inds]
    tmu = torch.tensor(list(mu), dtype=torch.float)
    B1 = torch.tensor(np.asarray(basis), dtype=torch.float)
    B2 = torch.tensor(np.asarray(basis2), dtype=torch.float)
    y0 = B1.T @ torch.tensor(np.asarray(w0).ravel(), dtype=torch.float)
    idx = np.concatenate((aug, n_cells + aug))
    V, Vbar = B1[idx, :], B2[idx, :]

    def decode32(x):
        return V @ x + Vbar @ net(torch.cat((x, tmu)))

    def decode(x):
        with torch.no_grad():
            return decode32(x)

    jacfwd = torch.func.jacfwd(decode32)
    red = np.zeros((y0.shape[0], num_steps + 1))
    red[:, 0] = y0.numpy()
    yp = y0.clone()
    wp = decode(y0).numpy().astype(np.float64)
    history = []
    for i in range(num_steps):
        y, resnorms = gauss_newton_manifold_ecsw(
            lambda x: residual_ecsw(x, grid_x, grid_y, dt, wp, mu, jdx_s, jdy_s, sample_inds, aug),
            lambda x: jacobian_ecsw(x, dt, jdx_s, jdy_s, eye2, sample_inds, aug),
            yp, decode, lambda x: jacfwd(x).detach(), sample_weights)
        history.append(resnorms)
        red[:, i + 1] = y.numpy()
        wp = decode(y).numpy().astype(np.float64)
        yp = y.clone()
    return red, history


def ecsw_training_matrix(snaps, prev_snaps, basis, grid_x, grid_y, dt, mu):
    """``compute_ECSW_training_matrix_2D`` (hypernet2D.py:2795-2816) with the exact Jacobian:
    ``C[i*k:(i+1)*k, cell] = r_u[cell]*(JV)_u[cell, :] + r_v[cell]*(JV)_v[cell, :]``."""
    n2, n_snaps = snaps.shape
    n_cells = n2 // 2
    k = basis.shape[1]
    C = np.zeros((k * n_snaps, n_cells))
    jdx, jdy = lifted_operators(grid_x, grid_y)
    for i in range(n_snaps):
        r = residual(snaps[:, i], grid_x, grid_y, dt, prev_snaps[:, i], mu)
        W = jacobian(snaps[:, i], dt, jdx, jdy).dot(basis)
        C[i * k:(i + 1) * k, :] = (r[:n_cells, None] * W[:n_cells] + r[n_cells:, None] * W[n_cells:]).T
    return C


# --------------------------------------------------------------------------------------
# small helpers at the boundary
# --------------------------------------------------------------------------------------
def compute_error(rom_snaps, hdm_snaps):
    """Per-snapshot relative error and its mean (hypernet2D.py:3150-3155).
    (The reference normalises by the ROM norm; kept.)"""
    sq_hdm = np.sqrt(np.square(rom_snaps).sum(axis=0))
    sq_err = np.sqrt(np.square(rom_snaps - hdm_snaps).sum(axis=0))
    rel_err = sq_err / sq_hdm
    return rel_err, rel_err.mean()


def param_to_snap_fn(mu, snap_folder="param_snaps", suffix=".npy"):
    """Cache file name ``param_snaps/mu1_<repr>+mu2_<repr>.npy`` (hypernet2D.py:3157-3181)."""
    return snap_folder + "/" + "+".join(
        "mu{}_{}".format(i + 1, m) for i, m in enumerate(mu)) + suffix
