"""LSPG PROM on the fine grid with the grid rows dealt to N GPUs.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \\
        --master-port 29511 tools/bench_lspg_sharded.py [--grid 750 --modes 95 --steps 10]

Rank 0 computes a POD basis from FOM snapshots on its device and broadcasts it; every rank keeps
the basis rows its stencils read.  Prints one JSON line (rank 0): ms per time step (device time,
max over ranks), Gauss-Newton counts, a checksum of the reduced trajectory (equal across N up to
the rounding of the all-reduce) and the single-GPU LspgSolver beside it at N = 1.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=750)
    ap.add_argument("--modes", type=int, default=95)
    ap.add_argument("--train-steps", type=int, default=120)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--jacobian", default="reference")
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    from finitedifference_b200.discretisation import make_2D_grid
    from finitedifference_b200.fom import FomSolver
    from finitedifference_b200.rom import LspgSolver, ShardedLspgSolver

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    n, k, dt, mu = a.grid, a.modes, 0.05, [4.75, 0.02]
    N = n * n
    gx, gy = make_2D_grid(0, 100, 0, 100, n, n)
    V = torch.empty((2 * N, k), dtype=torch.float64, device=dev)
    if rank == 0:
        fom = FomSolver(gx, gy, dt, a.train_steps, [[4.25, 0.015], [5.5, 0.03]], device=dev)
        fom.set_initial(torch.ones(2 * N, dtype=torch.float64, device=dev))
        fom.run()
        S = torch.cat([fom.snapshots_device(b, layout=0) for b in range(2)], dim=0)
        evals, evecs = torch.linalg.eigh(S @ S.T)
        order = torch.argsort(evals, descending=True)[:k]
        V.copy_(S.T @ (evecs[:, order] / torch.sqrt(evals[order])))
        del fom, S
    if world > 1:
        dist.broadcast(V, 0)
    Vh = V.cpu().numpy()
    del V
    torch.cuda.empty_cache()
    w0 = np.ones(2 * N)
    solver = ShardedLspgSolver(gx, gy, dt, mu, Vh, device=dev, jacobian=a.jacobian)
    solver.run(w0, 2)                                         # warm-up
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    snaps, red, hist = solver.run(w0, a.steps)
    e1.record()
    e1.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    line = {"stage": "lspg_sharded", "n_gpus": world, "grid": n, "modes": k, "steps": a.steps, "jacobian": a.jacobian,
            "ms_per_step": float(ms.item()) / a.steps, "gn_residual_evals": sum(len(h) for h in hist),
            "gn_its": [len(h) for h in hist], "rows_per_rank": solver.row_range[1] - solver.row_range[0],
            "basis_fraction_held": float(solver.state_cells.size) / N,
            "red_checksum": float(red.abs().sum().item()), "red_last": red[-1, :3].tolist()}
    if world == 1:
        single = LspgSolver(gx, gy, dt, mu, Vh, device=dev, jacobian=a.jacobian)
        single.run(w0, 2)
        torch.cuda.synchronize()
        e0.record()
        s1, r1, h1 = single.run(w0, a.steps)
        e1.record()
        e1.synchronize()
        line["single_gpu_solver_ms_per_step"] = e0.elapsed_time(e1) / a.steps
        line["max_rel_diff_vs_single"] = float(((red - r1).norm(dim=1) / r1.norm(dim=1)).max().item())
        line["gn_its_equal"] = [len(h) for h in h1] == line["gn_its"]
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
