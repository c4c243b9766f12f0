import sys, json, torch
sys.path.insert(0, '/root/repo')
from tools import records as R
dev = torch.device("cuda", 0)
for T in (20, 60):
    for k in ("chunk", "tile"):
        r = R.slab_record(dev, 0, 1, None, T=T, kernel=k, check=512 if T == 20 else 0)
        print(k, T, r["ms_per_trajectory"], r["cell_updates_per_s"] / 1e9, r["steps_redone"], r["newton_its_mean"], r["bit_exact_vs_single_gpu"], flush=True)
