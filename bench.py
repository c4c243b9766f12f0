#!/usr/bin/env python
"""bench.py -- FOM cell-updates/s of the parameter-sweep snapshot generation on the fine grid.

Workload (BASELINE.json configs[2], the configuration the metric is quoted on; configs[1] is
its one-sample case): the implicit full-order 2D Burgers stepper on the 750 x 750 grid, dt=0.05,
500 time steps, w0 = 1, for a lattice of mu = (mu1, mu2) in [4.25, 5.5] x [0.015, 0.03];
every GPU advances `--mu-per-gpu` samples together (weak scaling, no communication).

One bench "step" = one pass of the hot path over one batch: all `mu-per-gpu` trajectories,
all 500 time steps, Newton to 1e-12 each.  value = cell-updates/s over all GPUs with the
initial state resident in HBM and the snapshots left in HBM; e2e = the same through the
drop-in API with host buffers (w0 from pinned host memory in, (2N, T+1) snapshot arrays out to
pinned host memory) inside the timed region.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GRID_N = 750
DT = 0.05
NUM_STEPS = 500
MU1_RANGE = (4.25, 5.5)
MU2_RANGE = (0.015, 0.03)
NCU_DRAM_BYTES_PER_CELL_UPDATE = 43.1     # ncu --set full, profiles/r01_fom_full_ncu.md
FP64_INSTR_PER_CELL_UPDATE_K4 = 218       # DFMA/DMUL/DADD issued per cell-update with 4 Newton solves (ncu source page)
FP64_PEAK_TFLOPS = 34.09                  # measured DFMA peak, profiles/r01_fp64_microbench.txt
METRIC = "fom_cell_updates_per_s"
UNIT = "cell-updates/s"


def mu_lattice(count, rank=0, world=1):
    """`count` samples per rank (weak scaling) on a lattice of the training range, split contiguously."""
    from finitedifference_b200 import sharding
    return sharding.shard_mus(sharding.mu_lattice(count * world, MU1_RANGE, MU2_RANGE), rank, world)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *exc):
        if self.proc:
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, val in zip(names, r[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        busy = sorted(sm)[len(sm) // 2:]          # upper half = samples under load
        return {"sm_mhz": float(np.median(busy)), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}


def algorithmic_bytes(its, cells):
    """48 B per cell-sweep (read x_k, read wp, write x_{k+1}); 32 B for the first sweep of a
    step where x_0 == wp is read once.  its[m, t] = residual evaluations (solves + 1) of sample m
    at step t, as newton_raphson counts them."""
    its = np.asarray(its, dtype=np.int64)
    return int((cells * (32 + 48 * (its - 1))).sum())


def cpu_baseline(sample_steps=1, n=GRID_N):
    """The oracle (a NumPy/SciPy port of the reference: sparse Jacobian + SuperLU) on the host."""
    from oracle import burgers_oracle as orc
    gx, gy = orc.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    t0 = time.perf_counter()
    _, hist = orc.fom_trajectory(gx, gy, w0, DT, sample_steps, [4.75, 0.02], return_info=True)
    el = time.perf_counter() - t0
    return {"value": n * n * sample_steps / el, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": f"first {sample_steps} time step(s) of one mu on the {n}x{n} grid, "
                      f"{sum(len(h) for h in hist)} residual evaluations, {el:.1f} s "
                      "(SuperLU spsolve is single-threaded)",
            "seconds": el}


def run_reference_arm(args, rank, world):
    """--impl reference: the reference algorithm's CPU path (oracle port) on the host cores.
    Each step is a bounded sample of the workload: one time step of one mu on the fine grid."""
    if rank != 0:
        return
    times = []
    n = args.grid
    for i in range(args.warmup + args.steps):
        b = cpu_baseline(sample_steps=1, n=n)
        if i >= args.warmup:
            times.append(b["seconds"])
    total = sum(times)
    value = n * n * len(times) / total
    b["value"] = value
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"FOM {n}x{n}, dt={DT}: one time step of one mu per bench step "
                                   "(bounded sample of the 500-step sweep)", "grid": n,
                       "cpu_count": os.cpu_count()},
            "cpu_baseline": {k: b[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mu-per-gpu", type=int, default=int(os.environ.get("BFD_MU_PER_GPU", "24")))
    ap.add_argument("--chunk-steps", type=int, default=int(os.environ.get("BFD_CHUNK", "0")),
                    help="time steps fused per launch (0 = library default)")
    ap.add_argument("--non-strict", action="store_true", help="do not redo over-speculated steps")
    ap.add_argument("--num-steps", type=int, default=NUM_STEPS, help="time steps per trajectory (500 = the named config)")
    ap.add_argument("--grid", type=int, default=GRID_N)
    ap.add_argument("--e2e-mu", type=int, default=24, help="samples per end-to-end step (<= mu-per-gpu)")
    ap.add_argument("--e2e-group", type=int, default=2, help="samples a sweep worker advances together")
    ap.add_argument("--e2e-workers", type=int, default=3, help="sweep worker threads / streams")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--kernel", default="chunk", choices=["chunk", "tile"],
                    help="FOM sweep kernel: fom_chunk_kernel (default) or fom_tile_kernel (warp per strip and time step)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    # stdout carries exactly one JSON line: whatever libraries print there (NCCL's version banner
    # ignores NCCL_DEBUG_FILE) is sent to stderr, the line is written to the saved descriptor
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import ctypes as C
    import torch
    import torch.distributed as dist
    from finitedifference_b200 import _lib
    from finitedifference_b200.fom import FomSolver
    from finitedifference_b200.discretisation import make_2D_grid

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # keep NCCL's banner / debug lines off stdout: one JSON line only
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.get_lib()
    n, T = args.grid, args.num_steps
    cells = n * n
    gx, gy = make_2D_grid(0, 100, 0, 100, n, n)

    # size the batch to the HBM actually free
    skw = dict(chunk_steps=args.chunk_steps, strict=not args.non_strict, kernel=args.kernel)
    one = _lib.fom_desc(n, n, 1, T, chunk_steps=args.chunk_steps)
    per_mu = lib.bfd_fom_workspace_bytes(C.byref(one))
    free, _total = torch.cuda.mem_get_info(dev)
    B = max(1, min(args.mu_per_gpu, int(0.90 * free // per_mu)))
    mus = mu_lattice(B, rank, world)
    w0_dev = torch.ones(2 * cells, dtype=torch.float64, device=dev)

    solver = FomSolver(gx, gy, DT, T, mus, device=dev, **skw)
    ev = lambda: torch.cuda.Event(enable_timing=True)

    def one_step():
        e0, e1, e2 = ev(), ev(), ev()
        e0.record(); solver.set_initial(w0_dev); e1.record(); solver.run(); e2.record()
        e2.synchronize()
        return e0.elapsed_time(e2), e1.elapsed_time(e2)

    for _ in range(args.warmup):
        one_step()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    launches0 = lib.bfd_launch_count()
    with ClockSampler(local_rank) as clocks:
        wall0 = time.perf_counter()
        step_ms, run_ms = zip(*[one_step() for _ in range(args.steps)])
        torch.cuda.synchronize()
        wall = time.perf_counter() - wall0
    launches = lib.bfd_launch_count() - launches0
    total_ms = float(sum(step_ms))
    if world > 1:
        t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
        dist.barrier()
    its, rel = solver.diagnostics()
    sweeps = solver.launches
    redone = solver.steps_redone
    value = world * B * cells * T * args.steps / (total_ms * 1e-3)

    # roofline of the sweep kernel (the only kernel inside run()): algorithmic bytes / device time
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    abytes = algorithmic_bytes(its, cells)
    achieved = abytes / (float(np.mean(run_ms)) * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": NCU_DRAM_BYTES_PER_CELL_UPDATE * cells * T * B / max(sweeps, 1),
                "kernel": "fom_chunk_kernel",
                "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6.65 TB/s",
                "algorithmic_bytes_per_step": abytes, "launches_per_step": sweeps,
                "avg_launch_us": 1e3 * float(np.mean(run_ms)) / max(sweeps, 1),
                "note": "algorithmic bytes = (48*solves + 32) B per cell-update (SURVEY 8d: one HBM pass per "
                        "Newton sweep). The kernel fuses the sweeps of a time step in registers, so its real DRAM "
                        "traffic is lower (traffic, from ncu) and the binding unit is the FP64 pipe (fp64)",
                "traffic_source": "profiles/r01_fom_full_ncu.md: dram read+write = 43.1 B per cell-update "
                                  "(ncu --set full of this command), scaled to this run's cell-updates per launch",
                "fp64": {"instr_per_cell_update": FP64_INSTR_PER_CELL_UPDATE_K4,
                         "achieved_tflops": 2e-12 * FP64_INSTR_PER_CELL_UPDATE_K4 * value / world,
                         "peak_tflops": FP64_PEAK_TFLOPS, "frac": 2e-12 * FP64_INSTR_PER_CELL_UPDATE_K4 * value
                         / world / FP64_PEAK_TFLOPS,
                         "source": "profiles/r01_fp64_microbench.txt (DFMA peak measured on this pool) and the "
                                   "SASS count of FP64 instructions per cell-update at 4 solves"},
                "steps_redone": redone}

    # end to end through the public sweep API: host w0 in, every (2N, T+1) snapshot array out to pinned
    # host memory (finitedifference_b200/sweep.py: groups of samples on worker streams, the copies of
    # one group overlapping the sweeps of the next; host buffers recycled through a pool)
    del solver
    torch.cuda.empty_cache()
    from finitedifference_b200.sweep import SnapshotSweep, bind_host_to_gpu_node
    numa = bind_host_to_gpu_node(dev)           # pinned buffers and worker threads next to this GPU's PCIe root
    Be = max(1, min(args.e2e_mu, B))
    mus_e = mus[:Be]
    w0_np = np.ones(2 * cells)
    # pinned host buffers: a pool per rank, sized to the host memory the ranks of this box share
    import psutil
    buf_bytes = 2 * cells * (T + 1) * 8
    pool = int(0.35 * psutil.virtual_memory().available / max(world, 1) / buf_bytes)
    pool = max(2, min(args.e2e_group * args.e2e_workers, pool))
    sweep = SnapshotSweep(gx, gy, DT, T, device=dev, group=args.e2e_group, workers=args.e2e_workers, pool=pool,
                          **skw)
    seen = {}

    def sink(i, mu, snaps):
        # the trajectory is on the host: first column is w0, last column is the final state
        seen[i] = (float(snaps[::4099, 0].sum()), float(snaps[::4099, -1].sum()))

    sweep.run(w0_np, mus_e[:args.e2e_group * args.e2e_workers], sink=sink)        # warm-up: buffers, solvers
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e2e_reps = max(1, min(args.steps, 2))
    t0 = time.perf_counter()
    for _ in range(e2e_reps):
        seen.clear()
        sweep.run(w0_np, mus_e, sink=sink)
        assert len(seen) == Be and all(np.isfinite(v[1]) and v[0] == len(w0_np[::4099]) for v in seen.values()), seen
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e = {"value": world * Be * cells * T * e2e_reps / e2e_s, "unit": UNIT,
           "h2d_bytes_per_step": int(w0_np.nbytes),
           "d2h_bytes_per_step": int(Be * 2 * cells * (T + 1) * 8),
           "mu_per_gpu": Be, "seconds_per_step": e2e_s / e2e_reps, "pinned_buffers_per_gpu": pool,
           "host_numa_node": numa,
           "d2h_GBps_per_gpu": Be * 2 * cells * (T + 1) * 8 * e2e_reps / e2e_s / 1e9,
           "api": f"SnapshotSweep(group={args.e2e_group}, workers={args.e2e_workers}).run(w0, mus, sink): every "
                  "(2N, T+1) trajectory copied to pinned host memory and handed to the sink"}
    del sweep

    if rank == 0:
        cb = None
        if not args.no_cpu_baseline and world == 1:       # the CPU arm is timed at N = 1 only
            cb = cpu_baseline(sample_steps=1, n=n)
            cb = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": f"FOM parameter sweep, {n}x{n} grid, dt={DT}, {T} steps, "
                                       f"{B} mu per GPU (lattice over mu1 {MU1_RANGE} x mu2 {MU2_RANGE}), "
                                       "Newton rel tol 1e-12",
                           "grid": n, "num_steps": T, "mu_per_gpu": B, "chunk_steps": args.chunk_steps or 50,
                           "strict": not args.non_strict,
                           "l2": f"working set {B * per_mu / 2**30:.1f} GiB of snapshots per GPU, far larger than L2; "
                                 "no flush needed", "newton_its_mean": float(its.mean()),
                           "max_relnorm": float(rel.max())},
                "mu_snapshots_per_s": world * B * T * args.steps / (total_ms * 1e-3),
                "mu_trajectories_per_s": world * B * args.steps / (total_ms * 1e-3),
                "roofline": roofline, "cpu_baseline": cb, "e2e": e2e, "gpu_launches": int(launches),
                "clocks": clocks.summary(), "wall_s_timed": wall}
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
