"""Host-side discretisation constants, computed with the reference's own expressions.

All arrays are float64 NumPy; the CUDA kernels receive them verbatim, so the coefficients
the GPU multiplies with are bit-identical to the ones the reference's sparse operators hold.
"""
from __future__ import annotations

import numpy as np


def make_2D_grid(x_low, x_up, y_low, y_up, num_cells_x, num_cells_y):
    """Cell-edge coordinates; same as the reference's ``make_2D_grid`` (hypernet2D.py:2501-2507)."""
    grid_x = np.linspace(x_low, x_up, num_cells_x + 1)
    grid_y = np.linspace(y_low, y_up, num_cells_y + 1)
    return grid_x, grid_y


def stencil_coefficients(grid_x, grid_y, dt):
    """``cx, cxl, cy, cyd``: the entries of ``0.5*dt*Dx`` / ``0.5*dt*Dy`` (make_ddx, :2486-2492).

    ``cx[c] = 0.5*dt*(1/dx[c])`` is the diagonal, ``cxl[c] = 0.5*dt*(1/dx[c-1])`` minus the
    sub-diagonal (0 at c == 0: a missing neighbour contributes nothing).
    """
    grid_x = np.asarray(grid_x, dtype=np.float64)
    grid_y = np.asarray(grid_y, dtype=np.float64)
    dx = grid_x[1:] - grid_x[:-1]
    dy = grid_y[1:] - grid_y[:-1]
    h = 0.5 * dt
    cx = h * (1.0 / dx)
    cxl = np.zeros_like(cx); cxl[1:] = h * (1.0 / dx[:-1])
    cy = h * (1.0 / dy)
    cyd = np.zeros_like(cy); cyd[1:] = h * (1.0 / dy[:-1])
    return cx, cxl, cy, cyd


def parameter_terms(grid_x, grid_y, dt, mus):
    """Per-sample source and inlet terms: ``src[m, c]`` (:2625) and ``lbc[m, r]`` (:2628-2631).

    The reference writes ``lbc[:, 0] = 0.5*dt*mu1**2/dx`` -- the *array* dx broadcast down the
    first column, i.e. ``dx[r]`` (needs nx == ny).  For nx != ny the reduced-mesh residual's
    ``dx[0]`` (:2665) is used.
    """
    grid_x = np.asarray(grid_x, dtype=np.float64)
    grid_y = np.asarray(grid_y, dtype=np.float64)
    mus = np.atleast_2d(np.asarray(mus, dtype=np.float64))
    dx = grid_x[1:] - grid_x[:-1]
    ny = grid_y.size - 1
    xc = (grid_x[1:] + grid_x[:-1]) / 2
    src = np.stack([dt * 0.02 * np.exp(mu[1] * xc) for mu in mus])
    den = dx if dx.size == ny else np.full(ny, dx[0])
    lbc = np.stack([0.5 * dt * mu[0] ** 2 / den for mu in mus])
    return np.ascontiguousarray(src), np.ascontiguousarray(lbc)


def is_uniform(cy, cyd, rtol=1e-12):
    """True when the y coefficients are one value to ``rtol`` (``make_2D_grid`` grids are:
    linspace spacing varies by a few ulp), so the kernels may hold them in a register."""
    cy = np.asarray(cy); cyd = np.asarray(cyd)
    if cy.size < 2:
        return True
    ref = cy[1]
    return bool(np.all(np.abs(cy - ref) <= rtol * abs(ref)) and np.all(np.abs(cyd[1:] - ref) <= rtol * abs(ref)))
