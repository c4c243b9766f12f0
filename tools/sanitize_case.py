"""One small invocation of every kernel family, for compute-sanitizer (one tool per run):

    compute-sanitizer --tool racecheck|synccheck|memcheck python tools/sanitize_case.py

FOM sweep (both kernels: two strips, batch of two, a redo and a continuation pass), residual, gather and
grid-walk J V, normal equations (DMMA), q x q solve, V y, training-matrix block.  Results are checked
against the oracle so that a tool that perturbs timing cannot hide a wrong answer.  Slab runs are left
out on purpose: their kernels wait on one another across launches, which a tool that serialises
launches would turn into a hang."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    from finitedifference_b200 import hypernet2D as h, rom
    from oracle import burgers_oracle as orc
    n, steps, mu = 40, 3, [4.75, 0.02]
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    ref, hist = orc.fom_trajectory(gx, gy, w0, 0.05, steps, mu, return_info=True)
    for kernel in ("chunk", "tile"):
        for kw in ({}, dict(newton_hint=1, chunk_steps=2), dict(newton_hint=6)):
            snaps, info = h.inviscid_burgers_implicit2D(gx, gy, w0, 0.05, steps, mu, return_info=True, kernel=kernel, **kw)
            err = (np.linalg.norm(snaps - ref, axis=0) / np.linalg.norm(ref, axis=0)).max()
            assert err < 1e-10 and list(info["newton_its"]) == [len(x) for x in hist], (kernel, kw, err)
            print("fom", kernel, kw, "ok", err, flush=True)
    base = rom._RomBase()
    rng = np.random.default_rng(0)
    k = 37
    V = base._dev(rng.standard_normal((2 * n * n, k)))
    w = base._dev(1.0 + rng.random(2 * n * n))
    wp = base._dev(1.0 + rng.random(2 * n * n))
    for mode in ("exact", "reference"):
        mesh = rom.MeshTables(gx, gy, 0.05, mu, base.device, jacobian=mode)
        a, b = base.jv(mesh, w, V), base.jv(mesh, w, V, gather=True)
        assert torch.allclose(a, b, rtol=1e-13, atol=1e-13)
    mesh = rom.MeshTables(gx, gy, 0.05, mu, base.device)
    r, nr = base.residual(mesh, w, wp)
    assert abs(nr - np.linalg.norm(orc.residual(w.cpu().numpy(), gx, gy, 0.05, wp.cpu().numpy(), mu))) < 1e-10 * nr
    keep, rom.GRAM_KERNEL_MIN_ROWS = rom.GRAM_KERNEL_MIN_ROWS, 0
    JV = base.jv(mesh, w, V)
    G, bb = base.normal_equations(JV, r)
    rom.GRAM_KERNEL_MIN_ROWS = keep
    assert float((G - JV.T @ JV).abs().max()) < 1e-11 * float(G.abs().max())
    dy = base.gn_step(G, bb)
    assert float((G @ dy + bb).abs().max()) < 1e-9 * float(bb.abs().max())
    y = base._dev(rng.standard_normal(k))
    assert torch.allclose(base.basis_times(V, y), V @ y, rtol=1e-13, atol=1e-12)
    C = rom.ecsw_training_matrix(w.cpu().numpy()[:, None], wp.cpu().numpy()[:, None], V, gx, gy, 0.05, mu)
    Co = orc.ecsw_training_matrix(w.cpu().numpy()[:, None], wp.cpu().numpy()[:, None], V.cpu().numpy(), gx, gy, 0.05, mu)
    assert np.allclose(C, Co, rtol=1e-10, atol=1e-12)
    print("rom kernels ok", flush=True)


if __name__ == "__main__":
    main()
