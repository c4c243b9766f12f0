"""Read-only streaming bandwidth on this GPU, for reading the roofline fractions of the kernels that
only read (V y, the normal equations): torch reductions over a large tensor, cuBLAS GEMV and
bfd_basis_times_y on 95- and 96-column bases (760- vs 768-byte rows).  One JSON line."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from finitedifference_b200 import rom  # noqa: E402


def timed(fn, reps=10):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / reps


base = rom._RomBase()
dev = base.device
out = {}
x = torch.ones(1 << 29, dtype=torch.float64, device=dev)              # 4 GiB
out["torch_sum_4GiB_GBps"] = x.numel() * 8 / timed(lambda: x.sum()) / 1e6
x2 = x[: 1 << 27]
out["torch_sum_1GiB_GBps"] = x2.numel() * 8 / timed(lambda: x2.sum()) / 1e6
dst = torch.empty_like(x2)
out["torch_copy_1GiB_read_plus_write_GBps"] = 2 * x2.numel() * 8 / timed(lambda: dst.copy_(x2)) / 1e6
del x, x2, dst
M = 2 * 750 * 750
for k in (95, 96):
    V = torch.randn((M, k), dtype=torch.float64, device=dev)
    y = torch.randn(k, dtype=torch.float64, device=dev)
    out[f"gemv_cublas_k{k}_GBps"] = M * k * 8 / timed(lambda: V @ y) / 1e6
    out[f"basis_times_y_k{k}_GBps"] = M * k * 8 / timed(lambda: base.basis_times(V, y)) / 1e6
print(json.dumps(out))
