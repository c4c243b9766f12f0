"""Reduced-order models on one B200: host side of the LSPG PROM and the ECSW/ECM HPROM paths.

Mirrors ``inviscid_burgers_implicit2D_LSPG`` (reference hypernet2D.py:133-200) with
``gauss_newton_LSPG`` (:1935-2005), ``inviscid_burgers_ecsw_fixed`` (:202-273) with
``gauss_newton_ECSW_2D`` (:2007-2044), ``generate_augmented_mesh`` (:2522-2542) and
``compute_ECSW_training_matrix_2D`` (:2795-2816).

The host builds integer index tables (which state cells each residual row reads) once per
trajectory; residual rows and the matrix-free ``J @ V`` rows are hand-written kernels
(csrc/rom.cu) behind ``bfd_rom_residual`` / ``bfd_rom_jv`` / ``bfd_ecsw_c_block``.  The small
dense algebra of a Gauss-Newton step -- ``V @ y``, the k x k normal equations
``(JV)^T (JV) dy = -(JV)^T f`` (the reference's ``lstsq``; cond(JV) ~ 1.5) -- are
plain library GEMM / Cholesky calls on the device (cuBLAS / cuSOLVER through torch, float64).
The Gauss-Newton stopping tests are evaluated on the host from FP64 norms exactly as the
reference orders them.
"""
from __future__ import annotations

import ctypes as C
import time

import numpy as np
import torch

from . import _lib
from .discretisation import stencil_coefficients
from .fom import _require_cuda


GRAM_KERNEL_MIN_ROWS = 65536
SPD_SOLVE_KERNEL = True      # False: cuSOLVER potrf + potrs through torch (comparison)
JV_ROWS_PER_WARP = 0     # grid rows a warp of bfd_rom_jv_grid owns: 0 = library default, 1, 2 or 4 (tuning)


def generate_augmented_mesh(grid_x, grid_y, sample_inds):
    """Sampled cells plus their (r-1, c) and (r, c-1) neighbours: sorted, unique, int64.

    Same set, order and dtype as the reference's ``generate_augmented_mesh``
    (hypernet2D.py:2522-2542), which collects them in a Python ``set`` and sorts.
    """
    nx = np.asarray(grid_x).size - 1
    s = np.asarray(sample_inds, dtype=np.int64).ravel()
    r, c = np.divmod(s, nx)
    return np.unique(np.concatenate((s, s[r >= 1] - nx, s[c >= 1] - 1))).astype(np.int64)


class MeshTables:
    """Index and coefficient tables of ``bfd_rom_mesh`` for a set of rows, on the device."""

    def __init__(self, grid_x, grid_y, dt, mu, device, rows=None, state_cells=None, weight=None,
                 jacobian="exact", inlet="auto"):
        gx = np.asarray(grid_x, dtype=np.float64)
        gy = np.asarray(grid_y, dtype=np.float64)
        nx, ny = gx.size - 1, gy.size - 1
        n_cells = nx * ny
        rows = np.arange(n_cells, dtype=np.int64) if rows is None else np.asarray(rows, dtype=np.int64)
        r, c = np.divmod(rows, nx)
        if state_cells is None:                       # state arrays hold the whole grid
            lookup = None
            ncells = n_cells
        else:                                         # state arrays hold `state_cells` (sorted) only
            state_cells = np.asarray(state_cells, dtype=np.int64)
            ncells = state_cells.size

        def local(cells, present):
            """Position of grid cells in the state arrays; -1 where absent."""
            out = np.full(cells.shape, -1, dtype=np.int64)
            if state_cells is None:
                out[present] = cells[present]
            else:
                pos = np.searchsorted(state_cells, cells[present])
                pos = np.minimum(pos, state_cells.size - 1)
                if not np.array_equal(state_cells[pos], cells[present]):
                    raise ValueError("a stencil cell is missing from the augmented sample")
                out[present] = pos
            return out

        everywhere = np.ones(rows.shape, dtype=bool)
        p = local(rows, everywhere)
        xl = local(rows - 1, c >= 1)
        yd = local(rows - nx, r >= 1)
        if jacobian == "exact":
            ja, jb = p.copy(), yd.copy()
        elif jacobian == "reference":
            # hypernet2D.py:167: kron(I, Dyec)[permT, :] -- rows permuted, columns not: row (r, c)
            # reads cells (c, r) and (c, r-1)
            if nx != ny:
                raise ValueError("the reference's LSPG operator (hypernet2D.py:167) needs nx == ny")
            ja = local(c * ny + r, everywhere)
            jb = local(c * ny + r - 1, r >= 1)
        else:
            raise ValueError(jacobian)
        cx, cxl, cy, cyd = stencil_coefficients(gx, gy, dt)
        dx = gx[1:] - gx[:-1]
        xc = (gx[1:] + gx[:-1]) / 2
        src = (dt * 0.02 * np.exp(mu[1] * xc))[c]
        if inlet == "auto":                                               # whose inlet term: the full-grid residual's
            inlet = "grid" if state_cells is None else "sampled"         # or the sampled-mesh residual's
        if inlet == "grid":
            den = dx[r] if nx == ny else np.full(rows.shape, dx[0])      # :2582 / :2630 (Q3)
        else:
            den = np.full(rows.shape, dx[0])                              # :2665
        lbc = np.where(c == 0, 0.5 * dt * mu[0] ** 2 / den, 0.0)
        self._upload(dict(p=p, xl=xl, yd=yd, ja=ja, jb=jb, cx=cx[c], cxl=cxl[c], cy=cy[r], cyd=cyd[r], src=src,
                          lbc=lbc), ncells, device, weight)
        if state_cells is None and weight is None and rows.size == n_cells and np.array_equal(p, np.arange(n_cells)):
            # the whole grid in row order: J V can walk the grid (bfd_rom_jv_grid) instead of gathering
            self.grid = (nx, ny, int(jacobian == "reference"))

    @classmethod
    def from_tables(cls, tables, ncells, device, weight=None):
        """Tables given directly (keys as in ``bfd_rom_mesh``), e.g. read off caller-supplied operators."""
        self = cls.__new__(cls)
        self._upload(tables, ncells, device, weight)
        return self

    grid = None          # (nx, ny, transposed) when the rows are the whole grid in row order

    def with_terms(self, src=None, lbc=None):
        """The same mesh with caller-supplied per-row source / inlet terms (the reference's manifold drivers pass
        precomputed ``lbc`` and ``src`` to ``inviscid_burgers_res2D_ecsw``, hypernet2D.py:652, :856, :1081):
        index and coefficient tables are shared, only the two term arrays are uploaded."""
        other = MeshTables.__new__(MeshTables)
        other.nrows, other.ncells, other.device, other.grid = self.nrows, self.ncells, self.device, self.grid
        other._t = list(self._t)
        for slot, a in ((9, src), (10, lbc)):
            if a is not None:
                a = np.ascontiguousarray(np.asarray(a, dtype=np.float64).ravel())
                if a.size != self.nrows:
                    raise ValueError(f"expected {self.nrows} per-row terms, got {a.size}")
                other._t[slot] = torch.from_numpy(a).to(self.device)
        wptr = other._t[11].data_ptr() if len(other._t) > 11 else None
        other.c = _lib.RomMesh(other.nrows, other.ncells, *[x.data_ptr() for x in other._t[:11]], wptr)
        return other

    def _upload(self, t, ncells, device, weight):
        self.nrows, self.ncells, self.device = int(np.asarray(t["p"]).size), int(ncells), device
        i32 = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.int32)).to(device)
        f64 = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(device)
        self._t = [i32(t[k]) for k in ("p", "xl", "yd", "ja", "jb")] + \
                  [f64(t[k]) for k in ("cx", "cxl", "cy", "cyd", "src", "lbc")]
        wptr = None
        if weight is not None:
            self._t.append(f64(weight))
            wptr = self._t[-1].data_ptr()
        self.c = _lib.RomMesh(self.nrows, self.ncells, *[x.data_ptr() for x in self._t[:11]], wptr)


def _check_finite(norm, what):
    """The Gauss-Newton step comes from the k x k normal equations; a rank-deficient or hopelessly
    ill-conditioned J V (fewer sampled rows than modes, dependent basis columns, zero weights) makes
    them singular.  The reference's ``lstsq`` returns a minimum-norm step there; here that case is an
    error, not a silent NaN trajectory."""
    if not np.isfinite(norm):
        raise _lib.BurgersError(f"{what}: non-finite residual norm -- the normal equations of the Gauss-Newton step "
                                "were singular (rank-deficient J V: fewer sampled rows than modes, dependent basis "
                                "columns or zero weights)")


class _RomBase:
    def __init__(self, device=None, _backend=None):
        # _backend = (library, torch device): the CPU test-suite passes the SIMT-emulated build of
        # the same kernels (tests/simt_emu); the product always runs the CUDA library
        if _backend is None:
            self.lib = _lib.get_lib()
            self.device = _require_cuda(device)
        else:
            self.lib, self.device = _backend
        self.res_time = self.jac_time = self.ls_time = 0.0

    def _stream(self):
        if self.device.type != "cuda":
            return None
        return torch.cuda.current_stream(self.device).cuda_stream

    def _sync(self):
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)

    def _dev(self, a):
        return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(self.device)

    def residual(self, mesh, w, wp):
        """(r, ||r||) with r = [rows u | rows v] (weighted when the mesh has weights)."""
        r = torch.empty(2 * mesh.nrows, dtype=torch.float64, device=self.device)
        n2 = torch.zeros(1, dtype=torch.float64, device=self.device)
        _lib.check(self.lib.bfd_rom_residual(C.byref(mesh.c), w.data_ptr(), wp.data_ptr(), r.data_ptr(),
                                             n2.data_ptr(), self._stream()), self.lib, "bfd_rom_residual")
        return r, float(torch.sqrt(n2).item())

    def jv(self, mesh, w, V, out=None, gather=False):
        """Rows of ``J(w) @ V``: the grid-walking kernel on a whole-grid mesh, the gather kernel on a
        sampled mesh (or with ``gather=True``)."""
        k = V.shape[1]
        JV = torch.empty((2 * mesh.nrows, k), dtype=torch.float64, device=self.device) if out is None else out
        if mesh.grid is not None and not gather:
            nx, ny, transposed = mesh.grid
            _lib.check(self.lib.bfd_rom_jv_grid(C.byref(mesh.c), nx, ny, transposed, JV_ROWS_PER_WARP, w.data_ptr(),
                                                V.data_ptr(), k, JV.data_ptr(), self._stream()),
                       self.lib, "bfd_rom_jv_grid")
        else:
            _lib.check(self.lib.bfd_rom_jv(C.byref(mesh.c), w.data_ptr(), V.data_ptr(), k, JV.data_ptr(),
                                           self._stream()), self.lib, "bfd_rom_jv")
        return JV

    def basis_times(self, V, y, out=None):
        """``w = V @ y`` (:1999, :2042): the streaming kernel ``bfd_basis_times_y`` for k <= 96."""
        M, k = V.shape
        if k > 96 or M == 0 or not V.is_contiguous():
            return torch.mv(V, y, out=out) if out is not None else V @ y
        w = torch.empty(M, dtype=torch.float64, device=self.device) if out is None else out
        _lib.check(self.lib.bfd_basis_times_y(V.data_ptr(), y.contiguous().data_ptr(), M, k, w.data_ptr(), self._stream()),
                   self.lib, "bfd_basis_times_y")
        return w

    def normal_equations(self, JV, f):
        """``(G, b) = (JV^T JV, JV^T f)``: the DMMA kernel ``bfd_gram`` for k <= 96, library GEMMs beyond."""
        M, k = JV.shape
        if k > 96 or M < GRAM_KERNEL_MIN_ROWS:      # small sampled meshes: one library GEMM beats partials + reduce
            return JV.T @ JV, JV.T @ f
        ws = getattr(self, "_gram_ws", None)
        nbytes = self.lib.bfd_gram_workspace_bytes(M, k)
        if ws is None or ws.numel() < nbytes:
            ws = self._gram_ws = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        G = torch.empty((k, k), dtype=torch.float64, device=self.device)
        b = torch.empty(k, dtype=torch.float64, device=self.device)
        _lib.check(self.lib.bfd_gram(JV.data_ptr(), f.data_ptr(), M, k, G.data_ptr(), b.data_ptr(), ws.data_ptr(),
                                     ws.numel(), self._stream()), self.lib, "bfd_gram")
        return G, b

    def gn_step(self, G, b):
        """``dy = -G^-1 b`` = ``lstsq(JV, -f)``: one single-CTA launch (``bfd_spd_solve``) for k <= 96,
        cuSOLVER Cholesky beyond."""
        k = G.shape[0]
        if k > 96 or not SPD_SOLVE_KERNEL:
            L, _ = torch.linalg.cholesky_ex(G)
            return -torch.cholesky_solve(b[:, None], L)[:, 0]
        dy = torch.empty(k, dtype=torch.float64, device=self.device)
        _lib.check(self.lib.bfd_spd_solve(G.data_ptr(), b.data_ptr(), k, -1.0, dy.data_ptr(), self._stream()),
                   self.lib, "bfd_spd_solve")
        return dy

    def gn_step_min_norm(self, G, b):
        """Minimum-norm ``dy`` of a singular system ``G dy = -b`` (what ``lstsq(JV, -f)`` returns for a
        rank-deficient ``JV``): eigen-decomposition of G, eigenvalues below ``k eps max`` dropped."""
        lam, Q = torch.linalg.eigh(G)
        keep = lam > lam.max() * G.shape[0] * np.finfo(np.float64).eps
        return -(Q[:, keep] @ ((Q[:, keep].T @ b) / lam[keep]))

    def gauss_newton(self, res_mesh, jac_mesh, V, y0, wp, max_its=20, relnorm_cutoff=1e-5, min_delta=0.1):
        """``gauss_newton_LSPG`` / ``gauss_newton_ECSW_2D`` (:1935-2005, :2007-2044).

        Returns ``(y, w, resnorms)``; the two stopping rules are the reference's, in its order.  A
        Cholesky step that produced a non-finite iterate (singular normal equations) is retaken as
        the minimum-norm step; if that is not finite either the solver raises.
        """
        y = y0.clone()
        w = self.basis_times(V, y)
        f, init_norm = self.residual(res_mesh, w, wp)
        _check_finite(init_norm, "gauss_newton")
        resnorms = []
        retry = None
        for it in range(max_its):
            t0 = time.perf_counter()
            if it == 0:
                resnorm = init_norm          # the reference evaluates the same residual twice (:1964, :1970)
            else:
                f_new, resnorm = self.residual(res_mesh, w, wp)
                if not np.isfinite(resnorm) and retry is not None:
                    y_prev, G, b = retry
                    y = y_prev + self.gn_step_min_norm(G, b)
                    w = self.basis_times(V, y)
                    f_new, resnorm = self.residual(res_mesh, w, wp)
                _check_finite(resnorm, "gauss_newton")
                f = f_new
            self.res_time += time.perf_counter() - t0
            resnorms.append(resnorm)
            if resnorm / init_norm < relnorm_cutoff:
                break
            if len(resnorms) > 1 and abs((resnorms[-2] - resnorms[-1]) / resnorms[-2]) < min_delta:
                break
            t0 = time.perf_counter()
            JV = self.jv(jac_mesh, w, V)
            self._sync()
            self.jac_time += time.perf_counter() - t0
            t0 = time.perf_counter()
            G, b = self.normal_equations(JV, f)            # normal equations of lstsq(JV, -f)
            retry = (y, G, b)
            y = y + self.gn_step(G, b)
            w = self.basis_times(V, y)
            self._sync()
            self.ls_time += time.perf_counter() - t0
        return y, w, resnorms


class LspgSolver(_RomBase):
    """LSPG PROM on span(basis): ``inviscid_burgers_implicit2D_LSPG`` (:133-200).

    ``jacobian="reference"`` reproduces the operator the reference actually builds at :167
    (parity); ``"exact"`` uses the true Jacobian.
    """

    def __init__(self, grid_x, grid_y, dt, mu, basis, device=None, jacobian="reference"):
        super().__init__(device)
        self.grid_x, self.grid_y, self.dt, self.mu = grid_x, grid_y, float(dt), list(mu)[:2]
        self.V = basis if isinstance(basis, torch.Tensor) else self._dev(basis)
        self.V = self.V.to(self.device, dtype=torch.float64).contiguous()
        self.res_mesh = MeshTables(grid_x, grid_y, dt, self.mu, self.device)
        self.jac_mesh = self.res_mesh if jacobian == "exact" else \
            MeshTables(grid_x, grid_y, dt, self.mu, self.device, jacobian="reference")

    def run(self, w0, num_steps):
        """Returns ``(snaps_dev (T+1, 2N), red_coords_dev (T+1, k), per-step resnorm lists)``."""
        V = self.V
        w0 = w0 if isinstance(w0, torch.Tensor) else self._dev(np.asarray(w0, dtype=np.float64).ravel())
        y = V.T @ w0                                        # :158
        w = self.basis_times(V, y)                          # :159
        snaps = torch.empty((num_steps + 1, V.shape[0]), dtype=torch.float64, device=self.device)
        red = torch.empty((num_steps + 1, V.shape[1]), dtype=torch.float64, device=self.device)
        snaps[0], red[0] = w, y
        history = []
        for i in range(num_steps):
            y, w, resnorms = self.gauss_newton(self.res_mesh, self.jac_mesh, V, y, snaps[i])
            history.append(resnorms)
            snaps[i + 1], red[i + 1] = w, y                 # :196: the stored snapshot is basis @ y
        return snaps, red, history


class ShardedLspgSolver(_RomBase):
    """LSPG PROM with the grid rows dealt to the ranks of a process group.

    Rank ``g`` owns a contiguous block of grid rows: it evaluates the residual rows and the
    ``J V`` rows of its cells and forms its share of the normal equations; one all-reduce of
    ``[(JV)^T (JV) | (JV)^T f]`` (k x (k+1) doubles, 72 KB at k = 95) and one of ``||f||^2`` per
    Gauss-Newton iteration are the only exchanges, the reduced state ``y`` is replicated and
    every rank takes the same stopping decisions from the same reduced numbers
    (``gauss_newton_LSPG``, hypernet2D.py:1935-2005).  A rank keeps only the basis rows of the
    state cells its stencils read: the *row slab* (its rows and the row below) and, for the :167
    operator, the *column slab* of the transposed cells -- about 2/world of ``V``; ``w = V y`` is formed
    locally on exactly those cells, so no halo exchange exists.  ``J V`` walks the slab
    (``bfd_rom_jv_grid_slab``), the normal equations run on the FP64 tensor cores (``bfd_gram``).
    """

    def __init__(self, grid_x, grid_y, dt, mu, basis, rank=None, world=None, group=None, device=None,
                 jacobian="reference", _backend=None):
        import torch.distributed as dist
        super().__init__(device, _backend)
        self.dist, self.group = dist, group
        active = dist.is_available() and dist.is_initialized()
        self.rank = (dist.get_rank(group) if active else 0) if rank is None else int(rank)
        self.world = (dist.get_world_size(group) if active else 1) if world is None else int(world)
        from .sharding import shard_range
        gx, gy = np.asarray(grid_x, dtype=np.float64), np.asarray(grid_y, dtype=np.float64)
        nx, ny = gx.size - 1, gy.size - 1
        self.nx, self.ny, self.n_cells, self.mu = nx, ny, nx * ny, list(mu)[:2]
        if jacobian not in ("exact", "reference"):
            raise ValueError(jacobian)
        if jacobian == "reference" and nx != ny:
            raise ValueError("the reference's LSPG operator (hypernet2D.py:167) needs nx == ny")
        self.transposed = int(jacobian == "reference")
        block = shard_range(ny, self.rank, self.world)
        r_lo, r_hi = block.start, block.stop
        self.row_range = (r_lo, r_hi)
        row0 = max(r_lo - 1, 0)
        self.rows = np.arange(r_lo * nx, r_hi * nx, dtype=np.int64)
        pcells = np.arange(row0 * nx, r_hi * nx, dtype=np.int64)                     # row slab
        acells = (np.arange(nx, dtype=np.int64)[:, None] * ny + np.arange(row0, r_hi, dtype=np.int64)[None, :]).ravel() \
            if self.transposed else np.empty(0, dtype=np.int64)                       # column slab
        self.state_cells = np.union1d(pcells, acells)
        basis = basis.cpu().numpy() if isinstance(basis, torch.Tensor) else np.asarray(basis, dtype=np.float64)
        self.k = basis.shape[1]
        pick = lambda cells: np.concatenate((cells, self.n_cells + cells))
        self.nP, self.off = pcells.size, (r_lo - row0) * nx                           # owned cells start at `off` in the row slab
        self.V_P = self._dev(basis[pick(pcells)])
        self.V_A = self._dev(basis[pick(acells)]) if self.transposed and self.rows.size else None
        if self.rows.size:
            self.mesh = MeshTables(gx, gy, dt, self.mu, self.device, rows=self.rows, state_cells=pcells, inlet="grid")

    def _own(self, t):
        """The owned cells' [u | v] entries of a row-slab vector / matrix."""
        return torch.cat((t[self.off:self.nP], t[self.nP + self.off:]))

    def _allreduce(self, t):
        if self.world > 1:
            self.dist.all_reduce(t, group=self.group)
        return t

    def _residual(self, w, wp):
        """(f_local, ||f|| over all ranks)."""
        n2 = torch.zeros(1, dtype=torch.float64, device=self.device)
        f = torch.empty(2 * self.rows.size, dtype=torch.float64, device=self.device)
        if self.rows.size:
            _lib.check(self.lib.bfd_rom_residual(C.byref(self.mesh.c), w.data_ptr(), wp.data_ptr(), f.data_ptr(),
                                                 n2.data_ptr(), self._stream()), self.lib, "bfd_rom_residual")
        return f, float(torch.sqrt(self._allreduce(n2)).item())

    def _jv(self, w, wA):
        JV = torch.empty((2 * self.rows.size, self.k), dtype=torch.float64, device=self.device)
        _lib.check(self.lib.bfd_rom_jv_grid_slab(C.byref(self.mesh.c), self.nx, self.ny, self.transposed, self.row_range[0],
                                                 self.row_range[1], w.data_ptr(), self.V_P.data_ptr(),
                                                 wA.data_ptr() if self.transposed else None,
                                                 self.V_A.data_ptr() if self.transposed else None, self.k,
                                                 JV.data_ptr(), self._stream()), self.lib, "bfd_rom_jv_grid_slab")
        return JV

    def _gauss_newton(self, y0, wp, max_its=20, relnorm_cutoff=1e-5, min_delta=0.1):
        k = self.k
        y = y0.clone()
        w = self.basis_times(self.V_P, y)
        f, init_norm = self._residual(w, wp)
        resnorms = []
        for it in range(max_its):
            if it > 0:
                f, resnorm = self._residual(w, wp)
            else:
                resnorm = init_norm          # the reference evaluates the same residual twice (:1964, :1970)
            resnorms.append(resnorm)
            if resnorm / init_norm < relnorm_cutoff:
                break
            if len(resnorms) > 1 and abs((resnorms[-2] - resnorms[-1]) / resnorms[-2]) < min_delta:
                break
            Gb = torch.zeros((k, k + 1), dtype=torch.float64, device=self.device)
            if self.rows.size:
                JV = self._jv(w, self.basis_times(self.V_A, y) if self.transposed else None)
                G, b = self.normal_equations(JV, f)
                Gb[:, :k] = G
                Gb[:, k] = b
            self._allreduce(Gb)
            y = y + self.gn_step(Gb[:, :k].contiguous(), Gb[:, k].contiguous())
            w = self.basis_times(self.V_P, y)
        return y, w, resnorms

    def run(self, w0, num_steps):
        """``w0``: the full (2N,) initial state (host).  Returns ``(snaps_own (T+1, 2*n_own) -- the
        owned cells' [u | v] --, red_coords (T+1, k), per-step resnorm lists)`` on the device."""
        w0 = np.asarray(w0.cpu() if isinstance(w0, torch.Tensor) else w0, dtype=np.float64).ravel()
        own = np.concatenate((self.rows, self.n_cells + self.rows))
        y = self._allreduce(self._own(self.V_P).T @ self._dev(w0[own]))     # :158, summed over the row blocks
        w = self.basis_times(self.V_P, y)                                  # :159 on the row slab
        snaps = torch.empty((num_steps + 1, 2 * self.rows.size), dtype=torch.float64, device=self.device)
        red = torch.empty((num_steps + 1, self.k), dtype=torch.float64, device=self.device)
        snaps[0], red[0] = self._own(w), y
        history = []
        for i in range(num_steps):
            y, w, resnorms = self._gauss_newton(y, w)
            history.append(resnorms)
            snaps[i + 1], red[i + 1] = self._own(w), y                     # :196: the stored snapshot is basis @ y
        return snaps, red, history

    def gather_snapshots(self, snaps_own):
        """Full ``(2N, T+1)`` NumPy array on every rank from the per-rank row blocks."""
        T1 = snaps_own.shape[0]
        parts = [None] * self.world
        mine = (self.rows, snaps_own.cpu().numpy())
        if self.world > 1:
            self.dist.all_gather_object(parts, mine, group=self.group)
        else:
            parts = [mine]
        full = np.empty((2 * self.n_cells, T1))
        for rows, block in parts:
            n = rows.size
            full[rows] = block[:, :n].T
            full[self.n_cells + rows] = block[:, n:].T
        return full


class EcswSolver(_RomBase):
    """ECSW/ECM HPROM on the sampled mesh: ``inviscid_burgers_ecsw_fixed`` (:202-273)."""

    def __init__(self, grid_x, grid_y, weights, dt, mu, basis, device=None):
        super().__init__(device)
        weights = np.asarray(weights, dtype=np.float64).ravel()
        self.sample_inds = np.where(weights != 0)[0].astype(np.int64)             # :225 (Q7: != 0)
        self.augmented_sample = generate_augmented_mesh(grid_x, grid_y, self.sample_inds)
        n_cells = weights.size
        self.sample_weights = np.hstack((weights, weights))[self.sample_inds]     # :239-240
        self.idx = np.concatenate((self.augmented_sample, n_cells + self.augmented_sample))   # :242
        self.mu = list(mu)[:2]
        self.mesh = MeshTables(grid_x, grid_y, dt, self.mu, self.device, rows=self.sample_inds,
                               state_cells=self.augmented_sample, weight=self.sample_weights)
        basis = np.asarray(basis, dtype=np.float64) if not isinstance(basis, torch.Tensor) else basis
        self.V = basis if isinstance(basis, torch.Tensor) else self._dev(basis)
        self.V = self.V.to(self.device, dtype=torch.float64).contiguous()
        self.V_red = self.V[torch.from_numpy(self.idx).to(self.device)].contiguous()   # :245
        self.graph_kernel_launches = 0      # this library's kernels replayed inside CUDA graphs

    def run(self, w0, num_steps, graphs=True, speculate=True):
        """Returns ``(red_coords_dev (T+1, k), per-step resnorm lists)``.

        ``graphs=True`` (default): a Gauss-Newton iteration is one CUDA-graph launch and one host
        look at ``||f||`` (the stopping rules stay on the host, in the reference's order);
        ``graphs=False`` issues the same operations one by one.  ``speculate`` (graphs only): the iterations
        of a time step are launched back to back and judged afterwards, see ``_run_graphs``."""
        w0 = w0 if isinstance(w0, torch.Tensor) else self._dev(np.asarray(w0, dtype=np.float64).ravel())
        y = self.V.T @ w0                                   # :213
        red = torch.empty((num_steps + 1, self.V.shape[1]), dtype=torch.float64, device=self.device)
        red[0] = y
        wp = self.V_red @ y                                 # == (basis @ y0)[idx]  (:214, :243)
        history = []
        if graphs and self.device.type == "cuda":
            return self._run_graphs(y, wp, red, num_steps, speculate=speculate)
        for i in range(num_steps):
            y, w, resnorms = self.gauss_newton(self.mesh, self.mesh, self.V_red, y, wp)
            history.append(resnorms)
            red[i + 1] = y
            wp = w                                          # :271
        return red, history

    # -- Gauss-Newton iterations as CUDA graphs -------------------------------------------------
    def _capture(self):
        """Static buffers and two graphs: ``g_res`` (f = weighted residual of w, ||f||^2 to pinned host
        memory) and ``g_iter`` (J V, normal equations, Cholesky solve, y += dy, w = V y, then g_res's
        work for the new w)."""
        dev, V = self.device, self.V_red
        k, n_state, n_rows = V.shape[1], V.shape[0], 2 * self.mesh.nrows
        f64 = dict(dtype=torch.float64, device=dev)
        self._y, self._w, self._wp = torch.zeros(k, **f64), torch.zeros(n_state, **f64), torch.zeros(n_state, **f64)
        self._f, self._n2, self._JV = torch.zeros(n_rows, **f64), torch.zeros(1, **f64), torch.zeros((n_rows, k), **f64)
        self._h_n2 = torch.zeros(1, dtype=torch.float64).pin_memory()
        # evaluation j of a time step leaves ||f||^2 in _norms[j] and the iterate it was taken at in _ys[j]:
        # several iterations can then be launched back to back and judged afterwards (_run_graphs)
        self._norms, self._ys = torch.zeros(self.MAX_ITS + 2, **f64), torch.zeros((self.MAX_ITS + 2, k), **f64)
        self._it = torch.zeros(1, dtype=torch.long, device=dev)
        self._h_norms = torch.zeros(self.MAX_ITS + 2, dtype=torch.float64).pin_memory()
        mesh, lib = self.mesh, self.lib

        def residual_part():
            self._n2.zero_()
            _lib.check(lib.bfd_rom_residual(C.byref(mesh.c), self._w.data_ptr(), self._wp.data_ptr(), self._f.data_ptr(),
                                            self._n2.data_ptr(), self._stream()), lib, "bfd_rom_residual")
            self._h_n2.copy_(self._n2, non_blocking=True)
            self._norms.index_copy_(0, self._it, self._n2)
            self._ys.index_copy_(0, self._it, self._y[None, :])
            self._it.add_(1)

        def iteration():
            _lib.check(lib.bfd_rom_jv(C.byref(mesh.c), self._w.data_ptr(), V.data_ptr(), k, self._JV.data_ptr(),
                                      self._stream()), lib, "bfd_rom_jv")
            G, b = self.normal_equations(self._JV, self._f)
            self._y.add_(self.gn_step(G, b))                # y += lstsq(JV, -f)
            self.basis_times(V, self._y, out=self._w)
            residual_part()

        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):                       # warm-up outside capture (library handles, workspaces)
            self._w.copy_(V @ self._y)
            self._wp.copy_(self._w)
            residual_part()
            iteration()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self._g_res, self._g_iter = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._g_res):
            self._it.zero_()                                # evaluation 0 of a time step
            residual_part()
        with torch.cuda.graph(self._g_iter, pool=self._g_res.pool()):
            iteration()

    MAX_ITS = 20            # gauss_newton_ECSW_2D's max_its (:2008)

    def _run_graphs(self, y, wp, red, num_steps, max_its=20, relnorm_cutoff=1e-5, min_delta=0.1, speculate=True,
                    first_spec=2, adapt=True):
        """Time stepping with the Gauss-Newton iterations replayed as CUDA graphs.  ``speculate``: the
        iterations the previous time step needed are launched back to back after the first evaluation and
        judged afterwards from the norms they left behind -- one host synchronisation per time step instead
        of one per iteration; the stopping rules are applied to the same norms in the same order (:2022-2025),
        and when they would have stopped earlier the iterate of that evaluation is restored."""
        if max_its > self.MAX_ITS:
            raise ValueError(f"max_its <= {self.MAX_ITS}")

        def stop(r):
            return r[-1] / r[0] < relnorm_cutoff or (len(r) > 1 and abs((r[-2] - r[-1]) / r[-2]) < min_delta)

        with torch.cuda.device(self.device):
            if not hasattr(self, "_g_iter"):
                self._capture()
            stream = torch.cuda.current_stream(self.device)
            self._y.copy_(y)
            self._wp.copy_(wp)
            self._w.copy_(wp)                               # w of the first iterate: V_red @ y0 == wp
            history = []
            spec = first_spec if speculate else 0           # updates launched before the first look
            for i in range(num_steps):
                t0 = time.perf_counter()
                self._g_res.replay()
                launched = min(spec, max_its - 1)
                for _ in range(launched):
                    self._g_iter.replay()
                self._h_norms.copy_(self._norms, non_blocking=True)
                stream.synchronize()
                self.ls_time += time.perf_counter() - t0
                self.graph_kernel_launches += 1 + 2 * launched
                norms = np.sqrt(self._h_norms.numpy()[:launched + 1].copy())
                _check_finite(float(norms[0]), "EcswSolver")
                resnorms, done = [float(norms[0])], None
                for j in range(launched + 1):               # judge the evaluations in order
                    if j > 0:
                        resnorms.append(float(norms[j]))
                        _check_finite(resnorms[-1], "EcswSolver")
                    if stop(resnorms):
                        done = j
                        break
                if done is not None and done < launched:    # over-speculated: back to the iterate of evaluation `done`
                    self._y.copy_(self._ys[done])
                    self.basis_times(self.V_red, self._y, out=self._w)
                elif done is None:                          # more iterations, one look each (as without speculation)
                    for _ in range(max_its - 1 - launched):
                        t0 = time.perf_counter()
                        self._g_iter.replay()
                        stream.synchronize()
                        self.ls_time += time.perf_counter() - t0
                        self.graph_kernel_launches += 2
                        resnorms.append(float(np.sqrt(self._h_n2[0].item())))
                        _check_finite(resnorms[-1], "EcswSolver")
                        if stop(resnorms):
                            break
                    else:
                        # max_its evaluations without a stop: the reference still takes its last step (:2024-2043)
                        self._g_iter.replay()
                        stream.synchronize()
                history.append(resnorms)
                if speculate and adapt:                     # what this step needed is what the next one gets
                    spec = max(1, min(len(resnorms) - 1, 6))
                red[i + 1].copy_(self._y)
                self._wp.copy_(self._w)                     # :271
        return red, history


class EcswBatch:
    """ECSW HPROM trajectories for several test parameters on one GPU (replicas only).

    One HPROM step is a handful of tiny kernels and a host look at the norm, i.e. latency: the GPU
    is mostly idle.  Every parameter gets its own ``EcswSolver`` (sharing the basis in HBM), stream and
    captured graphs; one host thread launches the next Gauss-Newton iteration of every parameter
    that still iterates, then looks at their norms, so the iterations of different parameters
    overlap on the device.  Inside the graphs the normal equations are formed by ``bfd_gram`` (own
    workspace per solver): graphs that replay concurrently must not share a library workspace."""

    def __init__(self, grid_x, grid_y, weights, dt, mus, basis, device=None):
        global GRAM_KERNEL_MIN_ROWS
        base = _RomBase(device)
        self.device = dev = base.device
        self.V = basis.to(dev, dtype=torch.float64).contiguous() if isinstance(basis, torch.Tensor) else base._dev(basis)
        self.solvers, self.streams, self.events = [], [], []
        keep, GRAM_KERNEL_MIN_ROWS = GRAM_KERNEL_MIN_ROWS, 0
        try:
            with torch.cuda.device(dev):
                for mu in mus:                          # graphs are captured one at a time, here
                    solver = EcswSolver(grid_x, grid_y, weights, dt, mu, self.V, device=dev)
                    solver._capture()
                    self.solvers.append(solver)
                    self.streams.append(torch.cuda.Stream(device=dev))
                    self.events.append(torch.cuda.Event())
                torch.cuda.synchronize(dev)
        finally:
            GRAM_KERNEL_MIN_ROWS = keep

    def run(self, w0, num_steps, max_its=20, relnorm_cutoff=1e-5, min_delta=0.1):
        """``[(red_coords (k, T+1) ndarray, per-step resnorm lists), ...]`` in the order of ``mus``; each
        equals what ``inviscid_burgers_ecsw_fixed`` returns for that mu (same stopping rules)."""
        dev, S = self.device, self.solvers
        w0_dev = w0.to(dev) if isinstance(w0, torch.Tensor) else \
            torch.from_numpy(np.ascontiguousarray(np.asarray(w0, dtype=np.float64).ravel())).to(dev)
        n = len(S)
        reds, hists = [], [[] for _ in range(n)]

        def launch(i, graph):
            with torch.cuda.stream(self.streams[i]):
                graph.replay()
                self.events[i].record()

        def norm(i):
            self.events[i].synchronize()
            r = float(np.sqrt(S[i]._h_n2[0].item()))
            _check_finite(r, f"EcswBatch (parameter {i})")
            return r

        def stop(r, init):
            return r[-1] / init < relnorm_cutoff or (len(r) > 1 and abs((r[-2] - r[-1]) / r[-2]) < min_delta)

        with torch.cuda.device(dev):
            torch.cuda.synchronize(dev)
            for i, s in enumerate(S):
                with torch.cuda.stream(self.streams[i]):
                    y = s.V.T @ w0_dev
                    red = torch.empty((num_steps + 1, s.V.shape[1]), dtype=torch.float64, device=dev)
                    red[0] = y
                    s._y.copy_(y)
                    s._wp.copy_(s.V_red @ y)
                    s._w.copy_(s._wp)
                    reds.append(red)
            for step in range(num_steps):
                for i in range(n):
                    launch(i, S[i]._g_res)
                res = [[norm(i)] for i in range(n)]
                init = [r[0] for r in res]
                active = list(range(n))
                for it in range(max_its):
                    active = [i for i in active if not stop(res[i], init[i])]
                    if not active:
                        break
                    for i in active:
                        launch(i, S[i]._g_iter)
                    if it == max_its - 1:               # the reference's last step is not followed by an evaluation
                        break
                    for i in active:
                        res[i].append(norm(i))
                for i in range(n):
                    hists[i].append(res[i])
                    with torch.cuda.stream(self.streams[i]):
                        reds[i][step + 1].copy_(S[i]._y)
                        S[i]._wp.copy_(S[i]._w)
            torch.cuda.synchronize(dev)
        return [(np.ascontiguousarray(r.T.cpu().numpy()), h) for r, h in zip(reds, hists)]


def ecsw_batch(grid_x, grid_y, weights, w0, dt, num_steps, mus, basis, device=None, threads=None):
    """One-shot ``EcswBatch``: build, capture, run."""
    return EcswBatch(grid_x, grid_y, weights, dt, mus, basis, device=device).run(w0, num_steps)


_TRAINING_MESH = {}


def _training_mesh(grid_x, grid_y, dt, mu, dev):
    """Index tables of the full grid, kept for repeated calls with the same grid (one entry)."""
    key = (np.asarray(grid_x).tobytes(), np.asarray(grid_y).tobytes(), float(dt), float(mu[0]), float(mu[1]), str(dev))
    if key not in _TRAINING_MESH:
        _TRAINING_MESH.clear()
        _TRAINING_MESH[key] = MeshTables(grid_x, grid_y, dt, mu, dev)
    return _TRAINING_MESH[key]


def ecsw_training_matrix(snaps, prev_snaps, basis, grid_x, grid_y, dt, mu, device=None, out_device=False):
    """``compute_ECSW_training_matrix_2D`` (:2795-2816) with the exact Jacobian of ``get_ops``:
    ``C[i*k:(i+1)*k, cell] = r_u[cell]*(JV)_u[cell, :] + r_v[cell]*(JV)_v[cell, :]``."""
    base = _RomBase(device)
    dev = base.device
    mu = list(mu)[:2]
    to_dev = lambda a: a.to(dev, dtype=torch.float64) if isinstance(a, torch.Tensor) else base._dev(a)
    S, P, V = to_dev(snaps), to_dev(prev_snaps), to_dev(basis).contiguous()
    n2, n_snaps = S.shape
    n_cells, k = n2 // 2, V.shape[1]
    mesh = _training_mesh(grid_x, grid_y, dt, mu, dev)
    Cmat = torch.empty((k * n_snaps, n_cells), dtype=torch.float64, device=dev)
    for i in range(n_snaps):
        w, wp = S[:, i].contiguous(), P[:, i].contiguous()
        r, _ = base.residual(mesh, w, wp)
        blk = Cmat[i * k:(i + 1) * k]
        if mesh.grid is not None and not mesh.grid[2]:
            _lib.check(base.lib.bfd_ecsw_c_grid(C.byref(mesh.c), mesh.grid[0], mesh.grid[1], w.data_ptr(), r.data_ptr(),
                                                V.data_ptr(), k, blk.data_ptr(), n_cells, base._stream()),
                       base.lib, "bfd_ecsw_c_grid")
        else:
            _lib.check(base.lib.bfd_ecsw_c_fused(C.byref(mesh.c), w.data_ptr(), r.data_ptr(), V.data_ptr(), k,
                                                 blk.data_ptr(), n_cells, base._stream()),
                       base.lib, "bfd_ecsw_c_fused")
    return Cmat if out_device else Cmat.cpu().numpy()


class ManifoldEcswSolver(_RomBase):
    """Sampled-mesh HPROM with a caller-supplied DECODER.

    The reference's manifold ROMs -- ``inviscid_burgers_rnm2D_ecsw`` (:556-674, neural decoder),
    ``inviscid_burgers_pod_rbf_2D_global_ecsw`` (:973-1111) and ``inviscid_burgers_pod_gp_2D_ecsw``
    (:1113-1275) with their solvers ``gauss_newton_rnm_ecsw`` (:2167-2217), ``gauss_newton_pod_rbf_ecsw``
    (:2293), ``gauss_newton_pod_gp_ecsw`` (:2382) -- all iterate on reduced coordinates ``y`` with
    ``w = decode(y)`` on the augmented sample and the Gauss-Newton matrix ``J(w) @ d decode / d y``;
    residual and Jacobian are the same ``res2D_ecsw`` / ``exact_jac2D_ecsw`` the linear HPROM uses.
    Here the decoder and its Jacobian are callables on torch tensors living on the device:

        decode(y)      -> (2 n_aug,) float64 CUDA tensor  [u of augmented cells | v of augmented cells]
        decode_jac(y)  -> (2 n_aug, q) float64 CUDA tensor (row-major)

    and the sampled residual (``bfd_rom_residual``), ``J(w) @ (d decode / d y)`` (``bfd_rom_jv``: the
    decoder Jacobian takes the place of ``V_red``), the normal equations and the q x q solve are the
    kernels of ``EcswSolver``.  ``idx`` holds the state entries the decoder must produce (:599).
    """

    def __init__(self, grid_x, grid_y, weights, dt, mu, device=None):
        super().__init__(device)
        weights = np.asarray(weights, dtype=np.float64).ravel()
        self.sample_inds = np.where(weights != 0)[0].astype(np.int64)
        self.augmented_sample = generate_augmented_mesh(grid_x, grid_y, self.sample_inds)
        n_cells = weights.size
        self.sample_weights = np.concatenate((weights, weights))[self.sample_inds]           # :584
        self.idx = np.concatenate((self.augmented_sample, n_cells + self.augmented_sample))   # :599
        self.mu = list(mu)[:2]
        self.mesh = MeshTables(grid_x, grid_y, dt, self.mu, self.device, rows=self.sample_inds,
                               state_cells=self.augmented_sample, weight=self.sample_weights)

    def gauss_newton(self, y0, wp, decode, decode_jac, max_its=20, relnorm_cutoff=1e-5, min_delta=0.1,
                     update=None):
        """``gauss_newton_rnm_ecsw`` (:2167-2217): returns ``(y, w, resnorms)``.  ``update(y, dy)`` lets the
        caller apply the step in its own precision (the reference keeps ``y`` in float32)."""
        y = y0.clone()
        w = decode(y).to(torch.float64).contiguous()
        f, init_norm = self.residual(self.mesh, w, wp)
        _check_finite(init_norm, "ManifoldEcswSolver")
        resnorms = []
        for it in range(max_its):
            t0 = time.perf_counter()
            if it > 0:
                f, resnorm = self.residual(self.mesh, w, wp)
                _check_finite(resnorm, "ManifoldEcswSolver")
            else:
                resnorm = init_norm
            self.res_time += time.perf_counter() - t0
            resnorms.append(resnorm)
            if resnorm / init_norm < relnorm_cutoff:
                break
            if len(resnorms) > 1 and abs((resnorms[-2] - resnorms[-1]) / resnorms[-2]) < min_delta:
                break
            t0 = time.perf_counter()
            Vq = decode_jac(y).to(torch.float64).contiguous()
            JV = self.jv(self.mesh, w, Vq)
            self._sync()
            self.jac_time += time.perf_counter() - t0
            t0 = time.perf_counter()
            G, b = self.normal_equations(JV, f)
            dy = self.gn_step(G, b)
            y = update(y, dy) if update is not None else y + dy
            w = decode(y).to(torch.float64).contiguous()
            self._sync()
            self.ls_time += time.perf_counter() - t0
        return y, w, resnorms

    def run(self, y0, num_steps, decode, decode_jac, **gn_kw):
        """Time stepping of the manifold HPROM (:640-672).  Returns ``(red_coords (T+1, q) tensor, per-step
        resnorm lists)``."""
        y = y0.clone()
        red = [y.clone()]
        wp = decode(y).to(torch.float64).contiguous()
        history = []
        for _ in range(num_steps):
            y, w, resnorms = self.gauss_newton(y, wp, decode, decode_jac, **gn_kw)
            history.append(resnorms)
            red.append(y.clone())
            wp = w
        return torch.stack(red), history
