"""Kernel logic of the FOM path on CPU: the product's .cu sources compiled against the SIMT
emulator (tests/simt_emu) and driven through the same C ABI, compared with the oracle."""
import numpy as np
import pytest

from oracle import burgers_oracle as orc
from tests import emu_driver as E


@pytest.fixture(scope="module")
def lib():
    return E.load()


def rel_err(a, b):
    return (np.linalg.norm(a - b, axis=0) / np.linalg.norm(b, axis=0)).max()


@pytest.mark.parametrize("nx,ny,mus,kw", [
    (16, 16, [[4.75, 0.02]], {}),
    (40, 40, [[5.19, 0.026], [4.3, 0.03]], dict(chunk_steps=2)),   # two strips, batch of two
    (70, 33, [[5.0, 0.02]], dict(uniform_y=False)),               # non-square, ragged last strip, cy[] path
    (33, 20, [[5.0, 0.02]], dict(newton_hint=1, chunk_steps=1)),   # under-speculated: continuation passes
    (20, 20, [[4.4, 0.016]], dict(newton_hint=6)),                 # over-speculated: the step is redone
    (40, 72, [[4.75, 0.02]], {}),                                  # tall enough for the unmasked main phases (ny > 52)
    (40, 72, [[4.75, 0.02]], dict(newton_hint=1, chunk_steps=2)),  # ... of the continuation pass too
    (40, 40, [[5.19, 0.026], [4.3, 0.03]], dict(chunk_steps=5, kernel=1)),        # fom_tile_kernel: 3 time steps in flight per strip
    (70, 33, [[5.0, 0.02]], dict(uniform_y=False, kernel=1)),                     # ... ragged last strip, cy[] path
    (33, 20, [[5.0, 0.02]], dict(newton_hint=1, chunk_steps=4, kernel=1)),        # ... continuation passes
    (40, 72, [[4.75, 0.02]], dict(newton_hint=6, kernel=1)),                      # ... capped at 4 fused solves, redo, main phase
])
def test_emulated_fom_matches_oracle(lib, nx, ny, mus, kw):
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, nx, ny)
    w0 = np.ones(2 * nx * ny)
    stats = {}
    steps = 5 if kw.get("kernel") else 3
    snaps, its, rel = E.fom_run(lib, gx, gy, w0, 0.05, steps, mus, stats=stats, **kw)
    if "newton_hint" in kw:
        assert stats["redone"] >= 1
    for m, mu in enumerate(mus):
        ref, hist = orc.fom_trajectory(gx, gy, w0, 0.05, steps, mu, return_info=True)
        assert its[m].tolist() == [len(h) for h in hist]
        assert rel_err(snaps[m], ref) < 1e-13
        assert np.all(rel[m] < 1e-12)


def test_emulated_fom_polls_through_copies_when_the_status_memory_is_not_mapped(lib):
    """bfd_fom_run writes its status words into the caller's page-locked memory from the GPU when that memory is
    mapped (fom_status_kernel), and falls back to small device-to-host copies when it is not: same trajectory,
    same counts either way."""
    nx = ny = 16
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, nx, ny)
    w0 = np.ones(2 * nx * ny)
    mapped = E.fom_run(lib, gx, gy, w0, 0.05, 3, [[4.75, 0.02]], chunk_steps=1)
    lib.emu_set_host_unmapped(1)
    try:
        copied = E.fom_run(lib, gx, gy, w0, 0.05, 3, [[4.75, 0.02]], chunk_steps=1)
    finally:
        lib.emu_set_host_unmapped(0)
    assert np.array_equal(mapped[0], copied[0]) and np.array_equal(mapped[1], copied[1])


def test_emulated_fom_golden(lib, golden):
    g = golden("fom_n16.npz")
    n = int(g["n"])
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, n, n)
    snaps, its, _ = E.fom_run(lib, gx, gy, np.ones(2 * n * n), float(g["dt"]), 6, [list(g["mu"])])
    assert its[0].tolist() == list(g["newton_its"][:6])
    assert rel_err(snaps[0], g["snaps"][:, :7]) < 1e-13


def test_emulated_non_strict_keeps_converged_iterate(lib):
    """strict=False never redoes an over-speculated step: the iterate is the further-converged
    x_K, still a root of the step to the Newton tolerance and within 1e-10 of the reference."""
    n = 20
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    stats = {}
    snaps, its, rel = E.fom_run(lib, gx, gy, w0, 0.05, 3, [[4.75, 0.02]], newton_hint=5, strict=False,
                                stats=stats)
    ref = orc.fom_trajectory(gx, gy, w0, 0.05, 3, [4.75, 0.02])
    assert stats["redone"] == 0
    assert rel_err(snaps[0], ref) < 1e-12
    assert np.all(rel[0] < 1e-12)


def test_emulated_max_its_cap(lib):
    n = 16
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    snaps, its, rel = E.fom_run(lib, gx, gy, w0, 0.05, 2, [[4.75, 0.02]], max_its=2)
    assert its[0].tolist() == [2, 2]
    # reference semantics at the cap: two solves were applied (hypernet2D.py:1913-1933)
    wp = w0.copy()
    jdx, jdy = orc.lifted_operators(gx, gy)
    for i in range(2):
        x, _ = orc.newton_raphson(lambda x: orc.residual(x, gx, gy, 0.05, wp, [4.75, 0.02]),
                                  lambda x: orc.jacobian(x, 0.05, jdx, jdy), wp, max_its=2)
        assert np.linalg.norm(snaps[0][:, i + 1] - x) / np.linalg.norm(x) < 1e-13
        wp = x


def test_emulated_residual(lib):
    import ctypes as C
    from finitedifference_b200.discretisation import parameter_terms, stencil_coefficients
    nx, ny, dt, mu = 37, 21, 0.05, [4.9, 0.021]
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, nx, ny)
    rng = np.random.default_rng(5)
    w, wp = 1 + rng.random(2 * nx * ny), 1 + rng.random(2 * nx * ny)
    coefs = [np.ascontiguousarray(a) for a in stencil_coefficients(gx, gy, dt)]
    src, lbc = parameter_terms(gx, gy, dt, [mu])
    r = np.zeros_like(w)
    n2 = np.zeros(1)
    rc = lib.bfd_residual(nx, ny, *[E.ptr(a) for a in coefs], E.ptr(src), E.ptr(lbc), E.ptr(w), E.ptr(wp),
                          E.ptr(r), E.ptr(n2), None)
    assert rc == 0
    ref = orc.residual(w, gx, gy, dt, wp, mu)
    assert np.max(np.abs(r - ref)) < 1e-15
    assert abs(n2[0] - ref @ ref) < 1e-12 * (ref @ ref)
