"""Parity of the CUDA PROM / HPROM paths (drop-in API over the C ABI) with golden vectors generated
from the unmodified reference and with the oracle.  Tolerance (BASELINE.json north_star): reduced
states within 1e-8 relative, Gauss-Newton iteration counts equal, ECSW index sets bit-exact."""
import numpy as np
import pytest

from oracle import burgers_oracle as orc

pytestmark = pytest.mark.gpu
TOL = 1e-8


@pytest.fixture(scope="module")
def h():
    from finitedifference_b200 import hypernet2D
    return hypernet2D


def rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


def test_lspg_golden(h, golden):
    g = golden("lspg_n24.npz")
    n, steps = int(g["n"]), int(g["steps"])
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    basis = np.ascontiguousarray(g["basis"])
    snaps, times, info = h.inviscid_burgers_implicit2D_LSPG(gx, gy, w0, float(g["dt"]), steps, list(g["mu"]), basis,
                                                            return_info=True)
    assert snaps.shape == g["snaps"].shape and snaps.flags.c_contiguous and len(times) == 4
    assert info["gn_its"].tolist() == g["gn_its"].tolist()
    assert times[0] == int(g["num_its"])
    assert max(rel(snaps[:, i], g["snaps"][:, i]) for i in range(steps + 1)) < TOL
    assert np.array_equal(w0, np.ones(2 * n * n))


@pytest.mark.parametrize("mode", ["reference", "exact"])
def test_lspg_vs_oracle(h, mode):
    n, steps, mu = 40, 5, [5.19, 0.026]
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    fom = orc.fom_trajectory(gx, gy, w0, 0.05, 12, mu)
    basis = np.ascontiguousarray(np.linalg.svd(fom, full_matrices=False)[0][:, :7])
    snaps, times, info = h.inviscid_burgers_implicit2D_LSPG(gx, gy, w0, 0.05, steps, mu, basis, jacobian=mode,
                                                            return_info=True)
    ref, rtimes, red, hist = orc.lspg_trajectory(gx, gy, w0, 0.05, steps, mu, basis, jacobian_mode=mode,
                                                 return_info=True)
    assert info["gn_its"].tolist() == [len(x) for x in hist]
    assert max(rel(info["red_coords"][:, i], red[:, i]) for i in range(steps + 1)) < TOL
    assert max(rel(snaps[:, i], ref[:, i]) for i in range(steps + 1)) < TOL


def test_ecsw_golden(h, golden):
    g = golden("ecsw_n24.npz")
    n, steps = int(g["n"]), int(g["steps"])
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    red, times, info = h.inviscid_burgers_ecsw_fixed(gx, gy, g["weights"], w0, float(g["dt"]), steps, list(g["mu"]),
                                                     np.ascontiguousarray(g["basis"]), return_info=True)
    assert red.shape == g["red_coords"].shape and len(times) == 3
    assert info["sample_inds"].dtype == np.int64 and np.array_equal(info["sample_inds"], g["sample_inds"])
    assert info["augmented_sample"].dtype == np.int64
    assert np.array_equal(info["augmented_sample"], g["augmented_sample"])        # bit-exact index sets
    assert np.array_equal(h.generate_augmented_mesh(gx, gy, g["sample_inds"]), g["augmented_sample"])
    assert info["gn_its"].tolist() == g["gn_its"].tolist()
    assert max(rel(red[:, i], g["red_coords"][:, i]) for i in range(steps + 1)) < TOL


def test_ecsw_vs_oracle_with_boundary_weights(h):
    """Weights as run_HPROM_ecsw_joshua.py builds them: bc_w = 10 on every boundary cell, a
    sparse positive set inside."""
    n, steps, mu = 32, 6, [4.56, 0.019]
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    fom = orc.fom_trajectory(gx, gy, w0, 0.05, 10, mu)
    basis = np.ascontiguousarray(np.linalg.svd(fom, full_matrices=False)[0][:, :5])
    rng = np.random.default_rng(3)
    wts = np.zeros((n, n))
    wts[1:-1, 1:-1] = np.where(rng.random((n - 2, n - 2)) < 0.08, rng.uniform(0.8, 40.0, (n - 2, n - 2)), 0.0)
    wts[0, :] = wts[-1, :] = wts[:, 0] = wts[:, -1] = 10.0
    red, times, info = h.inviscid_burgers_ecsw_fixed(gx, gy, wts.ravel(), w0, 0.05, steps, mu, basis, return_info=True)
    ref, _, s_ref, a_ref, hist = orc.ecsw_trajectory(gx, gy, wts.ravel(), w0, 0.05, steps, mu, basis, return_info=True)
    assert np.array_equal(info["sample_inds"], s_ref) and np.array_equal(info["augmented_sample"], a_ref)
    assert info["gn_its"].tolist() == [len(x) for x in hist]
    assert max(rel(red[:, i], ref[:, i]) for i in range(steps + 1)) < TOL


def test_training_matrix_golden(h, golden):
    g = golden("ecswC_n12.npz")
    n, steps = int(g["n"]), int(g["steps"])
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    S = g["snaps"]
    Cm = h.compute_ECSW_training_matrix_2D(S[:, 3:steps:2], S[:, 0:steps - 3:2], g["basis"], None, None, gx, gy,
                                           float(g["dt"]), list(g["mu"]))
    assert Cm.shape == g["C"].shape
    assert np.allclose(Cm, g["C"], rtol=1e-11, atol=1e-15)


@pytest.mark.parametrize("n,k", [(97, 95), (64, 20), (130, 40)])
def test_grid_walk_jv_equals_gather_jv(n, k):
    """bfd_rom_jv_grid against bfd_rom_jv on the device, both operators, 2 and 4 rows per warp."""
    import torch
    from finitedifference_b200 import rom
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, n, n)
    rng = np.random.default_rng(n + k)
    base = rom._RomBase()
    w = base._dev(1.0 + rng.random(2 * n * n))
    V = base._dev(rng.standard_normal((2 * n * n, k)))
    for mode in ("exact", "reference"):
        mesh = rom.MeshTables(gx, gy, 0.05, [4.75, 0.02], base.device, jacobian=mode)
        assert mesh.grid == (n, n, int(mode == "reference"))
        want = base.jv(mesh, w, V, gather=True)
        for rows in (0, 2, 4):
            rom.JV_ROWS_PER_WARP = rows
            try:
                got = base.jv(mesh, w, V)
            finally:
                rom.JV_ROWS_PER_WARP = 0
            assert torch.allclose(got, want, rtol=1e-14, atol=1e-14), (mode, rows, float((got - want).abs().max()))
    Jo = orc.jacobian(w.cpu().numpy(), 0.05, *orc.lifted_operators(gx, gy)) if n <= 64 else None
    if Jo is not None:
        mesh = rom.MeshTables(gx, gy, 0.05, [4.75, 0.02], base.device)
        assert np.max(np.abs(base.jv(mesh, w, V).cpu().numpy() - Jo @ V.cpu().numpy())) < 1e-12


@pytest.mark.parametrize("mode", ["reference", "exact"])
def test_sharded_lspg_one_rank_equals_lspg(h, golden, mode):
    """ShardedLspgSolver (row blocks + all-reduce of the normal equations; here one rank, the
    multi-rank cases run on CPU in tests/test_lspg_sharded.py) against LspgSolver and the golden."""
    from finitedifference_b200.rom import LspgSolver, ShardedLspgSolver
    g = golden("lspg_n24.npz")
    n, steps = int(g["n"]), int(g["steps"])
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    sh = ShardedLspgSolver(gx, gy, float(g["dt"]), list(g["mu"]), g["basis"], rank=0, world=1, jacobian=mode)
    snaps_own, red, hist = sh.run(w0, steps)
    full = sh.gather_snapshots(snaps_own)
    s1, r1, h1 = LspgSolver(gx, gy, float(g["dt"]), list(g["mu"]), g["basis"], jacobian=mode).run(w0, steps)
    assert [len(x) for x in hist] == [len(x) for x in h1]
    assert max(rel(full[:, i], s1[i].cpu().numpy()) for i in range(steps + 1)) < 1e-12
    if mode == "reference":
        assert [len(x) for x in hist] == g["gn_its"].tolist()
        assert max(rel(full[:, i], g["snaps"][:, i]) for i in range(steps + 1)) < TOL


def over_start(solver, w0, steps):
    """(y, wp, red) as EcswSolver.run prepares them."""
    import torch
    w0d = solver._dev(np.asarray(w0, dtype=np.float64))
    y = solver.V.T @ w0d
    red = torch.empty((steps + 1, solver.V.shape[1]), dtype=torch.float64, device=solver.device)
    red[0] = y
    return y, solver.V_red @ y, red


def test_ecsw_graph_iterations_equal_eager(h, golden):
    """Gauss-Newton iterations replayed as CUDA graphs against the same operations issued one by one."""
    from finitedifference_b200.rom import EcswSolver
    g = golden("ecsw_n24.npz")
    n, steps = int(g["n"]), int(g["steps"])
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    args = (gx, gy, g["weights"], float(g["dt"]), list(g["mu"]), np.ascontiguousarray(g["basis"]))
    red_g, hist_g = EcswSolver(*args).run(w0, steps)
    red_e, hist_e = EcswSolver(*args).run(w0, steps, graphs=False)
    assert [len(x) for x in hist_g] == [len(x) for x in hist_e] == g["gn_its"].tolist()
    assert max(rel(red_g[i].cpu().numpy(), red_e[i].cpu().numpy()) for i in range(steps + 1)) < 1e-12
    solver = EcswSolver(*args)
    a, _ = solver.run(w0, 3)
    b, _ = solver.run(w0, 3)                       # captured graphs are reused by a second trajectory
    assert np.array_equal(a.cpu().numpy(), b.cpu().numpy())
    # iterations launched back to back and judged afterwards (default) against one look per iteration:
    # the same norms, the same counts, the same iterates -- also when the speculated count is too high
    # (forced here: six updates in flight where two are needed, so every step is rolled back)
    red_1, hist_1 = EcswSolver(*args).run(w0, steps, speculate=False)
    assert [len(x) for x in hist_1] == g["gn_its"].tolist()
    assert np.array_equal(red_g.cpu().numpy(), red_1.cpu().numpy())
    assert all(np.allclose(a, b, rtol=1e-13, atol=0) for a, b in zip(hist_g, hist_1))      # ||f||^2 is summed with atomics
    over = EcswSolver(*args)
    red_o, hist_o = over._run_graphs(*over_start(over, w0, steps), steps, speculate=True, first_spec=6, adapt=False)
    assert [len(x) for x in hist_o] == g["gn_its"].tolist()
    assert max(rel(red_o[i].cpu().numpy(), red_1[i].cpu().numpy()) for i in range(steps + 1)) < 1e-13


@pytest.mark.parametrize("M,k", [(200000, 95), (12345, 96), (999, 7), (70001, 40)])
def test_gram_dmma_equals_library_gemm(M, k, request):
    """bfd_gram (FP64 tensor cores, upper blocks only, fixed-order partial sums) against cuBLAS."""
    import torch
    from finitedifference_b200 import rom
    base = rom._RomBase()
    rom_min, rom.GRAM_KERNEL_MIN_ROWS = rom.GRAM_KERNEL_MIN_ROWS, 0          # the kernel at every size
    request.addfinalizer(lambda: setattr(rom, "GRAM_KERNEL_MIN_ROWS", rom_min))
    g = torch.Generator(device=base.device).manual_seed(M + k)
    X = torch.randn((M, k), dtype=torch.float64, device=base.device, generator=g)
    f = torch.randn(M, dtype=torch.float64, device=base.device, generator=g)
    G, b = base.normal_equations(X, f)
    Gr, br = X.T @ X, X.T @ f
    scale = float(Gr.abs().max())
    assert float((G - Gr).abs().max()) < 1e-12 * scale and torch.equal(G, G.T)
    assert float((b - br).abs().max()) < 1e-12 * scale
    G2, b2 = base.normal_equations(X, f)
    assert torch.equal(G, G2) and torch.equal(b, b2)          # deterministic


def test_ecsw_batch_equals_single_runs(h, golden):
    """Several test parameters iterating concurrently (own solver, stream and graphs each) reproduce the
    single-parameter drop-in call (to rounding: the batch forms the normal equations with bfd_gram)."""
    g = golden("ecsw_n24.npz")
    n, steps = int(g["n"]), int(g["steps"])
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    basis = np.ascontiguousarray(g["basis"])
    mus = [list(g["mu"]), [5.19, 0.026], [4.56, 0.019], [4.3, 0.016], [5.4, 0.029]]
    batch = h.inviscid_burgers_ecsw_fixed_batch(gx, gy, g["weights"], w0, float(g["dt"]), steps, mus, basis)
    assert len(batch) == len(mus)
    assert max(rel(batch[0][:, i], g["red_coords"][:, i]) for i in range(steps + 1)) < TOL
    for mu, red in zip(mus, batch):
        single, _ = h.inviscid_burgers_ecsw_fixed(gx, gy, g["weights"], w0, float(g["dt"]), steps, mu, basis)
        assert red.shape == single.shape and max(rel(red[:, i], single[:, i]) for i in range(steps + 1)) < 1e-12


def test_manifold_hprom_golden(h, golden):
    """inviscid_burgers_rnm2D_ecsw drop-in (decoder hook: ManifoldEcswSolver) against the reference's own
    run with a seeded MLP decoder.  The reference keeps the decoder and y in float32, so the tolerance
    here is float32 rounding through a few Gauss-Newton steps (1e-4), not the 1e-8 of the float64 HPROM."""
    from tests.test_oracle_golden import manifold_net
    g = golden("manifold_n24.npz")
    n, steps = int(g["n"]), int(g["steps"])
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    red, times, info = h.inviscid_burgers_rnm2D_ecsw(gx, gy, np.ones(2 * n * n), float(g["dt"]), steps, list(g["mu"]),
                                                     manifold_net(g), None, g["basis"], g["basis2"], g["weights"],
                                                     return_info=True)
    assert red.shape == g["red_coords"].shape and times[0] == int(g["num_its"])
    assert info["gn_its"].tolist() == g["gn_its"].tolist()
    assert max(rel(red[:, i], g["red_coords"][:, i]) for i in range(steps + 1)) < 1e-4


def test_manifold_hook_with_linear_decoder_equals_ecsw(h, golden):
    """The decoder hook with decode(y) = V_red y and a constant Jacobian is the linear HPROM: float64,
    reduced states within 1e-8 of the reference's inviscid_burgers_ecsw_fixed golden, counts equal."""
    import torch
    from finitedifference_b200.rom import ManifoldEcswSolver
    g = golden("ecsw_n24.npz")
    n, steps = int(g["n"]), int(g["steps"])
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    solver = ManifoldEcswSolver(gx, gy, g["weights"], float(g["dt"]), list(g["mu"]))
    V = torch.from_numpy(np.ascontiguousarray(g["basis"])).to(solver.device)
    Vr = V[torch.from_numpy(solver.idx).to(solver.device)].contiguous()
    y0 = V.T @ torch.ones(2 * n * n, dtype=torch.float64, device=solver.device)
    red, _, info = h.inviscid_burgers_manifold_ecsw(gx, gy, g["weights"], y0, float(g["dt"]), steps, list(g["mu"]),
                                                    lambda y: Vr @ y, lambda y: Vr, return_info=True)
    assert info["gn_its"].tolist() == g["gn_its"].tolist()
    assert np.array_equal(info["augmented_sample"], g["augmented_sample"])
    assert max(rel(red[:, i], g["red_coords"][:, i]) for i in range(steps + 1)) < TOL
