"""A few LSPG steps at 750^2, k = 95 on a random orthonormal basis: the shortest program that launches the
shipped J V (grid walk), Gram (FP64 tensor cores), solve and V y kernels in their production shapes.  Used under
`ncu -k regex:...` to capture them (profiles/r02_lspg_kernels_ncu.md); prints the ms per step it measured.

    python tools/lspg_probe.py [--steps 4]
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from finitedifference_b200 import hypernet2D as h
    from finitedifference_b200.rom import LspgSolver
    steps = int(sys.argv[sys.argv.index("--steps") + 1]) if "--steps" in sys.argv else 4
    dev = torch.device("cuda", 0)
    n, k = 750, 95
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    g = torch.Generator(device=dev).manual_seed(0)
    # smooth-ish columns (random combinations of a few low-frequency modes plus noise), then orthonormalised
    A = torch.randn((2 * n * n, k), generator=g, device=dev, dtype=torch.float64)
    A[:, 0] = 1.0
    V, _ = torch.linalg.qr(A)
    del A
    solver = LspgSolver(gx, gy, 0.05, [4.75, 0.02], V.contiguous(), device=dev, jacobian="reference")
    w0 = np.ones(2 * n * n)
    solver.run(w0, 1)                                                   # warm-up
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    _, _, hist = solver.run(w0, steps)
    t1.record()
    torch.cuda.synchronize()
    print(json.dumps({"steps": steps, "ms_per_step": t0.elapsed_time(t1) / steps,
                      "gn_residual_evals": [len(x) for x in hist]}))


if __name__ == "__main__":
    main()
