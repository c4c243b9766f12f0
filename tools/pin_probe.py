"""What returning a fresh (2N, T+1) host array costs at 750^2, T = 500 (4.5 GB): pinned, pageable,
pageable with huge pages requested.  The caller keeps the arrays (as run_prom.py does)."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from finitedifference_b200.discretisation import make_2D_grid  # noqa: E402
from finitedifference_b200.fom import FomSolver  # noqa: E402

n, T = 750, 500
gx, gy = make_2D_grid(0, 100, 0, 100, n, n)
s = FomSolver(gx, gy, 0.05, T, [[4.75, 0.02]])
s.set_initial(np.ones(2 * n * n))
s.run()
res = {"thp": open("/sys/kernel/mm/transparent_hugepage/enabled").read().strip()}
keep = []
ref = None
res["cpu_threads"] = torch.get_num_threads()
for name, kw in (("pageable_thp", dict()), ("staged_thp", dict(staged=True)), ("staged_thp_2", dict(staged=True))):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    a = s.snapshots(0, **kw)
    res[name] = time.perf_counter() - t0
    ref = a if ref is None else ref
    assert np.array_equal(a, ref), name
    keep.append(a)
print(res)
