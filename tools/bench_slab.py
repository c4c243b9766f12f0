"""BASELINE.json config 5: one large grid (default 8192 x 8192, single mu) over N GPUs by slab
decomposition, peer stores over NVLink for the one-cell-wide hand-off (finitedifference_b200/slab.py).

    python tools/bench_slab.py                                   # N = 1
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tools/bench_slab.py --gpus 2

Strong scaling: the grid is fixed, every rank holds 1/N of the columns.  Time = max over ranks of the
device time of run() (CUDA events), after a barrier.  --check first runs a smaller grid and compares
every rank's columns, bit for bit, with a single-GPU solve of the whole grid done on that same GPU.
Rank 0 prints one JSON line.
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
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--grid", type=int, default=8192)
    ap.add_argument("--num-steps", type=int, default=20)
    ap.add_argument("--steps", type=int, default=2, help="timed repetitions of the whole trajectory")
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--check", type=int, default=1024, help="grid size of the parity pre-check (0 = skip)")
    ap.add_argument("--chunk-steps", type=int, default=0)
    ap.add_argument("--dx", type=float, default=100.0 / 750.0, help="cell size (the fine configuration's)")
    a = ap.parse_args()

    import torch
    import torch.distributed as dist
    from finitedifference_b200 import _lib
    from finitedifference_b200.discretisation import make_2D_grid
    from finitedifference_b200.fom import FomSolver
    from finitedifference_b200.slab import FomSlab, connect_distributed

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # keep NCCL's banner / debug lines off stdout: one JSON line only
        dist.init_process_group("nccl", device_id=dev)
    barrier = (lambda: dist.barrier()) if world > 1 else (lambda: None)
    mu, dt = [4.75, 0.02], 0.05

    def make(n, T):
        # same cell size as the fine configuration (100/750): at dx = 100/8192 and dt = 0.05 the
        # reference's Newton iteration itself diverges (oracle), so the domain grows
        gx, gy = make_2D_grid(0, n * a.dx, 0, n * a.dx, n, n)
        s = FomSlab(gx, gy, dt, T, [mu], rank, world, device=dev, chunk_steps=a.chunk_steps)
        if world > 1:
            connect_distributed(s)
        else:
            s.connect()
        return gx, gy, s

    def trajectory(s):
        s.init()
        barrier()                                  # every rank's records are armed before anyone writes
        s.set_initial(torch.ones(2 * s.ny * s.nx_local, dtype=torch.float64, device=dev))
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s.run()
        e1.record()
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    check = None
    if a.check:
        n, T = a.check, 6
        gx, gy, s = make(n, T)
        trajectory(s)
        mine = s.snapshots_device(0)                                   # (T+1, 2, ny, nx_local)
        solo = FomSolver(gx, gy, dt, T, [mu], device=dev, chunk_steps=a.chunk_steps)
        solo.set_initial(torch.ones(2 * n * n, dtype=torch.float64, device=dev))
        solo.run()
        whole = solo.snapshots_device(0, layout=0).reshape(T + 1, 2, n, n)[:, :, :, s.c0:s.c1]
        same = bool(torch.equal(mine, whole))
        its, _ = s.diagnostics()
        its1, _ = solo.diagnostics()
        ok = torch.tensor([int(same and its.tolist() == its1.tolist())], device=dev)
        if world > 1:
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        check = {"grid": n, "steps": T, "bit_exact_vs_single_gpu": bool(ok.item()), "newton_its": its[0].tolist()}
        s.close()
        del solo, mine, whole
        torch.cuda.empty_cache()
        barrier()

    n, T = a.grid, a.num_steps
    gx, gy, s = make(n, T)
    lib = _lib.get_lib()
    for _ in range(a.warmup):
        trajectory(s)
    l0 = lib.bfd_launch_count()
    times = [trajectory(s) for _ in range(a.steps)]
    launches = lib.bfd_launch_count() - l0
    its, rel = s.diagnostics()
    ms = float(np.mean(times))
    if rank == 0:
        cells = n * n
        print(json.dumps({
            "metric": "fom_cell_updates_per_s", "value": cells * T / (ms * 1e-3), "unit": "cell-updates/s",
            "n_gpus": world, "scaling": "strong", "ms_per_trajectory": ms, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"single-mu FOM, {n}x{n} grid, {T} steps, slab decomposition over {world} GPU(s)",
                       "grid": n, "num_steps": T, "dx": a.dx, "strips_rank0": list(s.strips), "columns_rank0": [s.c0, s.c1]},
            "newton_its_mean": float(its.mean()), "newton_its": its[0].tolist(), "max_relnorm": float(rel.max()), "steps_redone": s.steps_redone,
            "launches_rank0": int(launches), "check": check,
            "halo_bytes_per_trajectory": int((world - 1) * T * n * (2 * float(its.mean()) + 1) * 16)}), flush=True)
    s.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
