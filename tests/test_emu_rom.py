"""ROM kernels (csrc/rom.cu) on CPU: the product's sources compiled against the SIMT emulator,
driven through the C ABI with the host-side index tables of finitedifference_b200/rom.py, and
compared with the oracle and with golden vectors generated from the unmodified reference."""
import ctypes as C

import numpy as np
import pytest
import torch

from finitedifference_b200 import _lib
from finitedifference_b200.rom import MeshTables, generate_augmented_mesh
from oracle import burgers_oracle as orc
from tests import emu_driver as E

CPU = torch.device("cpu")


@pytest.fixture(scope="module")
def lib():
    return E.load()


def run_residual(lib, mesh, w, wp):
    r = np.zeros(2 * mesh.nrows)
    n2 = np.zeros(1)
    assert lib.bfd_rom_residual(C.byref(mesh.c), E.ptr(w), E.ptr(wp), E.ptr(r), E.ptr(n2), None) == 0
    return r, n2[0]


def run_jv(lib, mesh, w, V):
    V = np.ascontiguousarray(V)
    JV = np.zeros((2 * mesh.nrows, V.shape[1]))
    assert lib.bfd_rom_jv(C.byref(mesh.c), E.ptr(w), E.ptr(V), V.shape[1], E.ptr(JV), None) == 0
    return JV


def test_full_mesh_residual_and_both_jacobians(lib, golden):
    g = golden("ops_n12.npz")
    n, dt, mu = int(g["n"]), float(g["dt"]), list(g["mu"])
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, n, n)
    w, wp = np.ascontiguousarray(g["w"]), np.ascontiguousarray(g["wp"])
    rng = np.random.default_rng(0)
    V = rng.standard_normal((2 * n * n, 5))
    exact = MeshTables(gx, gy, dt, mu, CPU)
    quirk = MeshTables(gx, gy, dt, mu, CPU, jacobian="reference")
    r, n2 = run_residual(lib, exact, w, wp)
    assert np.max(np.abs(r - g["r_std"])) < 1e-14
    assert abs(n2 - g["r_std"] @ g["r_std"]) < 1e-12 * n2
    assert np.max(np.abs(run_jv(lib, exact, w, V) - g["J_exact"] @ V)) < 1e-13      # get_ops operator
    assert np.max(np.abs(run_jv(lib, quirk, w, V) - g["J_lspg"] @ V)) < 1e-13       # hypernet2D.py:167


def run_jv_grid(lib, mesh, nx, ny, transposed, rows_per_warp, w, V):
    V = np.ascontiguousarray(V)
    JV = np.full((2 * mesh.nrows, V.shape[1]), np.nan)
    assert lib.bfd_rom_jv_grid(C.byref(mesh.c), nx, ny, int(transposed), rows_per_warp, E.ptr(w), E.ptr(V),
                               V.shape[1], E.ptr(JV), None) == 0
    return JV


@pytest.mark.parametrize("nx,ny,k", [(12, 12, 5), (37, 37, 40), (9, 9, 70), (45, 7, 33), (5, 50, 97), (20, 20, 95),
                                      (45, 7, 80), (3, 33, 96)])
def test_grid_walk_jv_equals_gather_jv(lib, nx, ny, k):
    """bfd_rom_jv_grid (walks the grid, neighbours in registers) against bfd_rom_jv (index tables):
    same arithmetic in the same order, for the exact operator and -- on square grids -- for the
    operator hypernet2D.py:167 builds; row counts that are no multiple of the rows a warp owns,
    mode counts that need 1, 2, 3 chunks of 32 lanes and more than one chunk group."""
    gx, gy = orc.make_2D_grid(0, 100, 0, 60, nx, ny)
    gx = gx + 0.3 * np.sin(gx)                                   # non-uniform spacing
    rng = np.random.default_rng(nx * 100 + k)
    w = 1.0 + rng.random(2 * nx * ny)
    V = rng.standard_normal((2 * nx * ny, k))
    modes = [("exact", 0)] + ([("reference", 1)] if nx == ny else [])
    for name, transposed in modes:
        mesh = MeshTables(gx, gy, 0.05, [4.3, 0.021], CPU, jacobian=name)
        want = run_jv(lib, mesh, w, V)
        for rows_per_warp in (0, 2, 4):
            got = run_jv_grid(lib, mesh, nx, ny, transposed, rows_per_warp, w, V)
            assert np.array_equal(got, want), (name, rows_per_warp, np.nanmax(np.abs(got - want)))
    mesh = MeshTables(gx, gy, 0.05, [4.3, 0.021], CPU)
    r = rng.standard_normal(2 * nx * ny)
    fused, walked = np.zeros((k, nx * ny + 3)), np.full((k, nx * ny + 3), np.nan)
    assert lib.bfd_ecsw_c_fused(C.byref(mesh.c), E.ptr(w), E.ptr(r), E.ptr(V), k, E.ptr(fused), nx * ny + 3, None) == 0
    assert lib.bfd_ecsw_c_grid(C.byref(mesh.c), nx, ny, E.ptr(w), E.ptr(r), E.ptr(V), k, E.ptr(walked), nx * ny + 3,
                               None) == 0
    assert np.array_equal(walked[:, :nx * ny], fused[:, :nx * ny]) and np.isnan(walked[:, nx * ny:]).all()
    if nx != ny:
        JV = np.zeros((2 * nx * ny, k))
        assert lib.bfd_rom_jv_grid(C.byref(mesh.c), nx, ny, 1, 0, E.ptr(w), E.ptr(V), k, E.ptr(JV), None) == -1
        assert lib.bfd_rom_jv_grid(C.byref(mesh.c), nx + 1, ny, 0, 0, E.ptr(w), E.ptr(V), k, E.ptr(JV), None) == -1


def test_sampled_mesh_matches_reference_rows(lib, golden):
    g = golden("ecsw_ops_n24.npz")
    n, dt, mu = int(g["n"]), float(g["dt"]), list(g["mu"])
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, n, n)
    weights = g["weights"]
    sample = np.where(weights != 0)[0]
    aug = generate_augmented_mesh(gx, gy, sample)
    assert aug.dtype == np.int64 and np.array_equal(aug, orc.augmented_mesh(gx, gy, sample))
    idx = np.concatenate((aug, n * n + aug))
    w, wp = np.ascontiguousarray(g["w"][idx]), np.ascontiguousarray(g["wp"][idx])
    mesh = MeshTables(gx, gy, dt, mu, CPU, rows=sample, state_cells=aug)
    r, _ = run_residual(lib, mesh, w, wp)
    assert np.max(np.abs(r - g["r_red"])) < 1e-14
    rng = np.random.default_rng(1)
    V = rng.standard_normal((2 * aug.size, 4))
    assert np.max(np.abs(run_jv(lib, mesh, w, V) - g["J_red"] @ V)) < 1e-13
    wts = np.hstack((weights, weights))[sample]
    meshw = MeshTables(gx, gy, dt, mu, CPU, rows=sample, state_cells=aug, weight=wts)
    rw, _ = run_residual(lib, meshw, w, wp)
    assert np.max(np.abs(rw - g["r_red"] * np.concatenate((wts, wts)))) < 1e-12
    JVw = run_jv(lib, meshw, w, V)
    assert np.max(np.abs(JVw - np.concatenate((wts, wts))[:, None] * (g["J_red"] @ V))) < 1e-11


def test_training_matrix_block(lib, golden):
    g = golden("ecswC_n12.npz")
    n, dt, mu = int(g["n"]), float(g["dt"]), list(g["mu"])
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, n, n)
    snaps, V = g["snaps"], np.ascontiguousarray(g["basis"])
    k, N = V.shape[1], n * n
    Cref = g["C"]
    n_snaps = Cref.shape[0] // k
    steps = int(g["steps"])
    cur, prev = snaps[:, 3:steps:2], snaps[:, 0:steps - 3:2]      # the pairing oracle/gen_golden.py used
    assert cur.shape[1] == n_snaps
    mesh = MeshTables(gx, gy, dt, mu, CPU)
    Cm = np.zeros_like(Cref)
    for i in range(n_snaps):
        w, wp = np.ascontiguousarray(cur[:, i]), np.ascontiguousarray(prev[:, i])
        r, _ = run_residual(lib, mesh, w, wp)
        JV = run_jv(lib, mesh, w, V)
        blk = np.zeros((k, N))
        assert lib.bfd_ecsw_c_block(N, k, E.ptr(r), E.ptr(JV), E.ptr(blk), N, None) == 0
        Cm[i * k:(i + 1) * k] = blk
        fused = np.zeros((k, N))           # the same block without the intermediate J V
        assert lib.bfd_ecsw_c_fused(C.byref(mesh.c), E.ptr(w), E.ptr(r), E.ptr(V), k, E.ptr(fused), N, None) == 0
        assert np.allclose(fused, blk, rtol=1e-13, atol=1e-16)
        walked = np.full((k, N), np.nan)   # ... and by walking the grid (bfd_ecsw_c_grid)
        assert lib.bfd_ecsw_c_grid(C.byref(mesh.c), n, n, E.ptr(w), E.ptr(r), E.ptr(V), k, E.ptr(walked), N, None) == 0
        assert np.array_equal(walked, fused)
    assert np.allclose(Cm, Cref, rtol=1e-11, atol=1e-15)
    assert np.allclose(Cm, orc.ecsw_training_matrix(cur, prev, V, gx, gy, dt, mu), rtol=1e-11, atol=1e-15)


def run_jac(lib, tables, ncells, w):
    from finitedifference_b200 import operators as ops
    mesh = MeshTables.from_tables(tables, ncells, CPU)
    vals = np.zeros((mesh.nrows, 12))
    assert lib.bfd_jac_csr_values(C.byref(mesh.c), E.ptr(np.ascontiguousarray(w)), E.ptr(vals), None) == 0
    return ops.assemble_csr(vals, tables, ncells)


def test_jacobian_csr_from_caller_operators(lib, golden):
    """inviscid_burgers_exact_jac2D as a matrix: tables read off the operators the caller passes
    (get_ops' and the rows-only-permuted one of hypernet2D.py:167), values from the kernel."""
    from finitedifference_b200 import operators as ops
    g = golden("ops_n12.npz")
    n, dt = int(g["n"]), float(g["dt"])
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, n, n)
    w = np.ascontiguousarray(g["w"])
    own = np.arange(n * n)
    dxec, dyec, jdx, jdy, eye = ops.get_ops(gx, gy)
    ox, oy = orc.lifted_operators(gx, gy)
    assert abs(jdx - ox).max() == 0 and abs(jdy - oy).max() == 0 and eye.shape == (2 * n * n,) * 2
    assert abs(dxec.tocsr() - orc.backward_difference(gx)).max() == 0 and dxec.format == "lil"
    J = run_jac(lib, ops.jacobian_tables(jdx, jdy, dt, own), n * n, w)
    assert J.shape == g["J_exact"].shape and np.max(np.abs(J.toarray() - g["J_exact"])) < 1e-14
    assert J.nnz == np.count_nonzero(g["J_exact"])
    qx, qy = orc.lifted_operators(gx, gy, "unpermuted")
    J = run_jac(lib, ops.jacobian_tables(qx, qy, dt, own), n * n, w)
    assert np.max(np.abs(J.toarray() - g["J_lspg"])) < 1e-14


def test_jacobian_csr_sampled_rows(lib, golden):
    from finitedifference_b200 import operators as ops
    g = golden("ecsw_ops_n24.npz")
    n, dt = int(g["n"]), float(g["dt"])
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, n, n)
    sample = np.where(g["weights"] != 0)[0]
    aug = generate_augmented_mesh(gx, gy, sample)
    jdx_s, jdy_s, eye2 = orc.ecsw_operators(gx, gy, sample, aug)
    rng = np.random.default_rng(5)
    w = 1.0 + rng.random(2 * aug.size)
    J = run_jac(lib, ops.jacobian_tables(jdx_s, jdy_s, dt, np.searchsorted(aug, sample)), aug.size, w)
    Jo = orc.jacobian_ecsw(w, dt, jdx_s, jdy_s, eye2, sample, aug)
    assert J.shape == Jo.shape and abs(J - Jo).max() < 1e-14
    idx = np.concatenate((aug, n * n + aug))
    J = run_jac(lib, ops.jacobian_tables(jdx_s, jdy_s, dt, np.searchsorted(aug, sample)), aug.size, g["w"][idx])
    assert np.max(np.abs(J.toarray() - g["J_red"])) < 1e-14                      # the reference's own matrix
    with pytest.raises(ValueError):
        ops.jacobian_tables(jdy_s, jdx_s, dt, np.searchsorted(aug, sample) + 1)


@pytest.mark.parametrize("M,k", [(1000, 95), (37, 5), (16, 96), (130, 33), (1, 1)])
def test_gram_dmma_normal_equations(lib, M, k):
    """bfd_gram (DMMA blocks on or above the diagonal, f as a 13th column block, partials summed
    in a fixed order) against NumPy: G = X^T X with both triangles filled, b = X^T f."""
    rng = np.random.default_rng(M * 100 + k)
    X = np.ascontiguousarray(rng.standard_normal((M, k)))
    f = rng.standard_normal(M)
    nbytes = lib.bfd_gram_workspace_bytes(M, k)
    assert nbytes > 0 and lib.bfd_gram_workspace_bytes(M, 97) == 0
    ws = np.zeros(nbytes // 8)
    G, b = np.full((k, k), np.nan), np.full(k, np.nan)
    assert lib.bfd_gram(E.ptr(X), E.ptr(f), M, k, E.ptr(G), E.ptr(b), E.ptr(ws), nbytes, None) == 0
    assert np.allclose(G, X.T @ X, rtol=1e-13, atol=1e-12) and np.array_equal(G, G.T)
    assert np.allclose(b, X.T @ f, rtol=1e-13, atol=1e-12)
    assert lib.bfd_gram(E.ptr(X), E.ptr(f), M, k, E.ptr(G), E.ptr(b), E.ptr(ws), nbytes - 8, None) == -3


@pytest.mark.parametrize("k", [1, 2, 7, 33, 95, 96])
def test_spd_solve_single_cta(lib, k):
    """bfd_spd_solve (packed Cholesky + triangular solves by one CTA) against numpy.linalg.solve."""
    rng = np.random.default_rng(k)
    A = rng.standard_normal((3 * k + 5, k))
    G = np.ascontiguousarray(A.T @ A)
    b = rng.standard_normal(k)
    x = np.full(k, np.nan)
    assert lib.bfd_spd_solve(E.ptr(G), E.ptr(b), k, -1.0, E.ptr(x), None) == 0
    want = -np.linalg.solve(G, b)
    assert np.allclose(x, want, rtol=1e-10, atol=1e-12 * np.abs(want).max())
    assert lib.bfd_spd_solve(E.ptr(G), E.ptr(b), 97, 1.0, E.ptr(x), None) == -1


@pytest.mark.parametrize("n,k,blocks", [(12, 5, [(0, 12)]), (13, 40, [(0, 5), (5, 6), (6, 13)]), (9, 70, [(0, 1), (1, 9)])])
def test_grid_walk_jv_on_row_slabs(lib, n, k, blocks):
    """bfd_rom_jv_grid_slab: the rows of a block of grid rows from slab-local state / basis arrays
    (row slab starting one row below; column slab of the transposed cells for hypernet2D.py:167)
    equal the same rows of the whole-grid call."""
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, n, n)
    gx = gx + 0.2 * np.cos(gx)
    rng = np.random.default_rng(n + k)
    N = n * n
    w = 1.0 + rng.random(2 * N)
    V = rng.standard_normal((2 * N, k))
    for name, transposed in (("exact", 0), ("reference", 1)):
        whole = run_jv_grid(lib, MeshTables(gx, gy, 0.05, [4.3, 0.021], CPU, jacobian=name), n, n, transposed, 0, w, V)
        for r_lo, r_hi in blocks:
            row0 = max(r_lo - 1, 0)
            own = np.arange(r_lo * n, r_hi * n)
            pcells = np.arange(row0 * n, r_hi * n)
            acells = (np.arange(n)[:, None] * n + np.arange(row0, r_hi)[None, :]).ravel()
            pick = lambda cells: np.concatenate((cells, N + cells))
            wP, VP = np.ascontiguousarray(w[pick(pcells)]), np.ascontiguousarray(V[pick(pcells)])
            wA, VA = np.ascontiguousarray(w[pick(acells)]), np.ascontiguousarray(V[pick(acells)])
            mesh = MeshTables(gx, gy, 0.05, [4.3, 0.021], CPU, rows=own, state_cells=pcells, inlet="grid")
            JV = np.full((2 * own.size, k), np.nan)
            assert lib.bfd_rom_jv_grid_slab(C.byref(mesh.c), n, n, transposed, r_lo, r_hi, E.ptr(wP), E.ptr(VP),
                                            E.ptr(wA) if transposed else None, E.ptr(VA) if transposed else None,
                                            k, E.ptr(JV), None) == 0
            assert np.array_equal(JV, whole[pick(own)]), (name, r_lo, r_hi)
    assert lib.bfd_rom_jv_grid_slab(C.byref(mesh.c), n, n, 1, r_lo, r_hi, E.ptr(wP), E.ptr(VP), None, None, k, E.ptr(JV),
                                    None) == -1


@pytest.mark.parametrize("M,k", [(1, 1), (7, 5), (64, 32), (101, 33), (250, 64), (333, 95), (40, 96)])
def test_basis_times_y(lib, M, k):
    """bfd_basis_times_y (four rows per warp, joint shuffle reduction) against NumPy."""
    rng = np.random.default_rng(M + k)
    V = np.ascontiguousarray(rng.standard_normal((M, k)))
    y = rng.standard_normal(k)
    w = np.full(M + 2, np.nan)
    assert lib.bfd_basis_times_y(E.ptr(V), E.ptr(y), M, k, E.ptr(w), None) == 0
    assert np.allclose(w[:M], V @ y, rtol=1e-13, atol=1e-13) and np.isnan(w[M:]).all()
    assert lib.bfd_basis_times_y(E.ptr(V), E.ptr(y), M, 97, E.ptr(w), None) == -1
