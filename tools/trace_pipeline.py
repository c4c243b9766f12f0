#!/usr/bin/env python
"""Print the strip pipeline of the fused FOM kernel: when each strip of one sample started and
finished each time step of the last launch (bfd_fom_trace).  GPU only; profiling aid.

    python tools/trace_pipeline.py [--grid 750] [--steps 26] [--batch 1] [--chunk 25]
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
    ap.add_argument("--steps", type=int, default=26)
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--chunk", type=int, default=25)
    ap.add_argument("--hint", type=int, default=0)
    args = ap.parse_args()
    import torch
    from finitedifference_b200.fom import FomSolver
    from finitedifference_b200.discretisation import make_2D_grid
    n = args.grid
    gx, gy = make_2D_grid(0, 100, 0, 100, n, n)
    mus = [[4.25 + 1.25 * i / max(args.batch - 1, 1), 0.02] for i in range(args.batch)]
    s = FomSolver(gx, gy, 0.05, args.steps, mus, chunk_steps=args.chunk, newton_hint=args.hint)
    w0 = torch.ones(2 * n * n, dtype=torch.float64, device="cuda")
    for rep in range(2):
        s.set_initial(w0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); s.run(); e1.record(); e1.synchronize()
    ns = (n + 31) // 32
    tr = torch.zeros((args.batch, args.chunk, ns, 3), dtype=torch.int64, device="cuda")
    rc = s.lib.bfd_fom_trace(C.byref(s.desc), s.ws.data_ptr(), tr.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    tr = tr.cpu().numpy()
    its, rel = s.diagnostics()
    print(f"run {e0.elapsed_time(e1):.2f} ms, launches {s.launches}, redone {s.steps_redone}, its {its[0][:8]}...")
    t = tr[0]
    base = t[..., 0][t[..., 0] > 0].min()
    rows = [i for i in range(args.chunk) if t[i, 0, 0] > 0]
    print("time step i: strip start (us) ... | duration of strip 0, mid, last (us)")
    for i in rows[:4] + rows[-2:]:
        st = (t[i, :, 0] - base) / 1e3
        du = (t[i, :, 1] - t[i, :, 0]) / 1e3
        print(i, "start", np.array2string(st[:: max(1, ns // 8)], precision=0),
              "dur", f"{du[0]:.0f} {du[ns // 2]:.0f} {du[-1]:.0f}", "end last", f"{(t[i, -1, 1] - base) / 1e3:.0f}",
              "misses/spins mid", int(t[i, ns // 2, 2]) >> 32, int(t[i, ns // 2, 2]) & 0xffffffff)


if __name__ == "__main__":
    main()
