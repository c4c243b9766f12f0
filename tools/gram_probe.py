"""Normal-equation kernels side by side for ncu: cuBLAS GEMM (JV^T JV), cuBLAS GEMV (JV^T f) and
bfd_gram (DMMA) on a 750^2, k = 95 sized problem."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from finitedifference_b200 import rom  # noqa: E402

M, k = 2 * 750 * 750, 95
base = rom._RomBase()
g = torch.Generator(device=base.device).manual_seed(0)
X = torch.randn((M, k), dtype=torch.float64, device=base.device, generator=g)
f = torch.randn(M, dtype=torch.float64, device=base.device, generator=g)
for _ in range(3):
    G = X.T @ X
    b = X.T @ f
    G2, b2 = base.normal_equations(X, f)
torch.cuda.synchronize()
print(float((G - G2).abs().max() / G.abs().max()))
