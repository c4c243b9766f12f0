"""bfd_spd_solve against cuSOLVER potrf + potrs at k = 95 (timing and difference); profiling aid."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from finitedifference_b200 import rom
from tools.records import timed
base = rom._RomBase()
for k in (95, 96, 40, 7):
    X = torch.randn((4000, k), dtype=torch.float64, device=base.device)
    G, b = X.T @ X, torch.randn(k, dtype=torch.float64, device=base.device)
    base.gn_step(G, b); torch.cuda.synchronize()          # first launch: module load, instruction cache
    dy, ms = timed(lambda: base.gn_step(G, b), 50)
    rom.SPD_SOLVE_KERNEL = False
    base.gn_step(G, b); torch.cuda.synchronize()
    ref, ms_lib = timed(lambda: base.gn_step(G, b), 50)
    rom.SPD_SOLVE_KERNEL = True
    print(f"k={k}: kernel {1e3 * ms:.1f} us, cuSOLVER {1e3 * ms_lib:.1f} us, max rel diff {float((dy - ref).abs().max() / ref.abs().max()):.2e}")
