#!/usr/bin/env python
"""Time-step pipeline of the FOM sweep kernels at the bench's shape (profiling aid, GPU only): for one
sample, when each strip started and finished each time step of the last launch (bfd_fom_trace), and how
long the strip waited for its left neighbour's start distance (third trace word, ns; tile kernel).

    python tools/trace_tile.py [--grid 750] [--steps 100] [--batch 24] [--chunk 50] [--kernel chunk|tile]
"""
import argparse
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=750)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--batch", type=int, default=24)
    ap.add_argument("--chunk", type=int, default=50)
    ap.add_argument("--kernel", default="chunk")
    ap.add_argument("--sample", type=int, default=0)
    args = ap.parse_args()
    import torch
    from finitedifference_b200.fom import FomSolver
    from finitedifference_b200.discretisation import make_2D_grid
    n = args.grid
    gx, gy = make_2D_grid(0, 100, 0, 100, n, n)
    mus = [[4.25 + 1.25 * i / max(args.batch - 1, 1), 0.02] for i in range(args.batch)]
    s = FomSolver(gx, gy, 0.05, args.steps, mus, chunk_steps=args.chunk, kernel=args.kernel)
    w0 = torch.ones(2 * n * n, dtype=torch.float64, device="cuda")
    for rep in range(2):
        s.set_initial(w0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); s.run(); e1.record(); e1.synchronize()
    ns = (n + 31) // 32
    tr = torch.zeros((args.batch, args.chunk, ns, 3), dtype=torch.int64, device="cuda")
    assert s.lib.bfd_fom_trace(C.byref(s.desc), s.ws.data_ptr(), tr.data_ptr(), torch.cuda.current_stream().cuda_stream) == 0
    tr = tr.cpu().numpy()
    ms = e0.elapsed_time(e1)
    print(f"{args.kernel}: run {ms:.2f} ms = {args.batch * n * n * args.steps / ms / 1e6:.2f} G cell-updates/s, "
          f"launches {s.launches}, redone {s.steps_redone}")
    t = tr[args.sample]
    live = t[..., 0] > 0
    base = t[..., 0][live].min()
    print("attempt: start (us) of strips 0, 1, 2, mid, last | duration (us) of the same | wait for the left strip (us)")
    sel = [0, 1, 2, ns // 2, ns - 1]
    atts = [i for i in range(args.chunk) if live[i, 0]]
    for i in atts[:8] + atts[-4:]:
        st = (t[i, sel, 0] - base) / 1e3
        du = (t[i, sel, 1] - t[i, sel, 0]) / 1e3
        wt = t[i, sel, 2] / 1e3
        print(f"{i:3d}", np.array2string(st, precision=0, suppress_small=True), "|", np.array2string(du, precision=0),
              "|", np.array2string(wt, precision=0))
    dur = (t[..., 1] - t[..., 0])[live] / 1e3
    wait = t[..., 2][live] / 1e3
    print(f"time-step duration per strip: median {np.median(dur):.0f} us, p10 {np.percentile(dur, 10):.0f}, p90 {np.percentile(dur, 90):.0f}; "
          f"wait at start: median {np.median(wait):.0f} us, mean {wait.mean():.0f} us")
    print(f"launch span of this sample: {(t[..., 1][live].max() - base) / 1e3:.0f} us for {len(atts)} attempts")


if __name__ == "__main__":
    main()
