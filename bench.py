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
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GRID_N = 750
DT = 0.05
NUM_STEPS = 500
MU1_RANGE = (4.25, 5.5)
MU2_RANGE = (0.015, 0.03)
METRIC = "fom_cell_updates_per_s"
UNIT = "cell-updates/s"
# Constants read off one ncu capture of the sweep kernel (profiles/r02_fom_chunk_ncu.md).  They describe the
# kernel SOURCE they were profiled on: the line carries `stale: true` when the sources' hash has moved on.
NCU_PROFILED_SHA = "c7c8e73c9e2256ab"             # sha256 of csrc/{fom.cu,wavefront.cuh,common.cuh} at profiling time, first 16 hex digits
NCU_DRAM_BYTES_PER_CELL_UPDATE = 43.0     # dram__bytes_read.sum + dram__bytes_write.sum per cell-update (r02 capture)
NCU_FP64_INSTR_PER_CELL_UPDATE = 227.3    # FP64-pipe warp instructions x 32 lanes / real cell-updates of the profiled launch, 4 solves (padding lanes included)
CPU_SECONDS_PER_CELL_UPDATE = 65e-6       # one core, first time step (5 solves) at 750^2: 36.5 s / 562 500 cells (r01)


def kernel_source_sha():
    import hashlib
    h = hashlib.sha256()
    for f in ("fom.cu", "wavefront.cuh", "common.cuh"):
        with open(os.path.join(ROOT, "finitedifference_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def mu_lattice(count, rank=0, world=1):
    """`count` samples per rank (weak scaling) on a lattice of the training range, split contiguously."""
    from finitedifference_b200 import sharding
    return sharding.shard_mus(sharding.mu_lattice(count * world, MU1_RANGE, MU2_RANGE), rank, world)


def algorithmic_bytes(its, cells):
    """48 B per cell-sweep (read x_k, read wp, write x_{k+1}); 32 B for the first sweep of a
    step where x_0 == wp is read once.  its[m, t] = residual evaluations (solves + 1) of sample m
    at step t, as newton_raphson counts them."""
    its = np.asarray(its, dtype=np.int64)
    return int((cells * (32 + 48 * (its - 1))).sum())


def workload_config(n, T, B, extra=None):
    cfg = {"workload": f"FOM parameter sweep, {n}x{n} grid, dt={DT}, {T} steps, {B} mu per GPU "
                       f"(lattice over mu1 {MU1_RANGE} x mu2 {MU2_RANGE}), Newton rel tol 1e-12",
           "grid": n, "num_steps": T, "mu_per_gpu": B}
    cfg.update(extra or {})
    return cfg


# ---- the reference's CPU path (oracle port: sparse Jacobian + SuperLU, as the reference does) -------------
def _cpu_band(job):
    """One time step of one mu on a band of `rows` grid rows of the n-column fine grid (same dx, dy, dt): the
    UNMODIFIED reference function when oracle/_ref/ has been made (oracle/make_ref.py), else the oracle port.
    Returns (seconds, residual evaluations, kind)."""
    n, rows, mu = job
    for v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[v] = "1"
    from oracle import burgers_oracle as orc
    from oracle import make_ref
    gx, gy = orc.make_2D_grid(0, 100, 0, 100.0 * rows / n, n, rows)
    w0 = np.ones(2 * n * rows)
    try:
        ref = make_ref.load()
    except Exception:                              # the baseline must not take the bench line down: fall back to the port
        ref = None
    if ref is not None:
        import contextlib
        import io
        counts = []
        newton = ref.newton_raphson

        def counted(*a, **k):                      # hypernet2D.py:1887: returns (x, resnorms)
            out = newton(*a, **k)
            counts.append(len(out[1]))
            return out
        ref.newton_raphson = counted
        try:
            with contextlib.redirect_stdout(io.StringIO()):
                t0 = time.perf_counter()
                ref.inviscid_burgers_implicit2D(gx, gy, w0, DT, 1, mu)          # hypernet2D.py:72
                dt = time.perf_counter() - t0
            return dt, sum(counts), "reference"
        except Exception:                          # reported as kind "port" by cpu_sample
            pass
        finally:
            ref.newton_raphson = newton
    t0 = time.perf_counter()
    _, hist = orc.fom_trajectory(gx, gy, w0, DT, 1, mu, return_info=True)
    return time.perf_counter() - t0, sum(len(h) for h in hist), "port"


def cpu_sample(n, rows, procs, pool=None):
    """`procs` parameter samples advanced one time step concurrently, one process per host core (the sweep
    is embarrassingly parallel over mu; SuperLU itself is single-threaded).  Returns the record."""
    import multiprocessing as mp
    mus = [[MU1_RANGE[0] + (MU1_RANGE[1] - MU1_RANGE[0]) * (i + 0.5) / procs, 0.02] for i in range(procs)]
    own = pool is None
    if own:
        from oracle import make_ref
        try:
            make_ref.load()             # imported once here; the forked workers inherit it (no import inside the timed region)
        except Exception:
            pass
        pool = mp.get_context("fork").Pool(procs)
    try:
        t0 = time.perf_counter()
        res = pool.map(_cpu_band, [(n, rows, m) for m in mus], chunksize=1)
        wall = time.perf_counter() - t0
    finally:
        if own:
            pool.close()
    evals = [r[1] for r in res]
    kind = "reference" if all(r[2] == "reference" for r in res) else "port"
    what = ("the UNMODIFIED reference function inviscid_burgers_implicit2D (oracle/_ref, made by oracle/make_ref.py)"
            if kind == "reference" else "the oracle port of the reference's path")
    return {"value": procs * n * rows / wall, "unit": UNIT, "cores": procs, "kind": kind, "seconds": wall,
            "sample": f"first time step of {procs} mu in parallel (one process per core, {os.cpu_count()} logical CPUs) on a "
                      f"{n} x {rows} band of the {n} x {n} grid (same dx, dy, dt), {min(evals)}-{max(evals)} residual "
                      f"evaluations each, {wall:.1f} s wall; {what}: sparse Jacobian + SuperLU spsolve",
            "solves_per_step": float(np.mean(evals)) - 1.0}


def cpu_rows(n, procs, seconds):
    """Band height that takes about `seconds` per step with `procs` processes sharing the host's memory system."""
    per_cell = CPU_SECONDS_PER_CELL_UPDATE * (1.0 + 0.04 * procs)
    return int(max(8, min(n, seconds / (n * per_cell))))


def host_cores():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def run_reference_arm(args, rank, world):
    """--impl reference: the reference's CPU path on all host cores -- the unmodified reference function when
    oracle/_ref/ exists (oracle/make_ref.py; `kind: "reference"`), else the oracle port (`kind: "port"`).  A bench
    step is a bounded sample of the workload, sized so that warm-up + timed steps end within ~150 s."""
    if rank != 0:
        return
    import multiprocessing as mp
    n, procs = args.grid, host_cores()
    total = args.warmup + args.steps
    rows = cpu_rows(n, procs, 150.0 / max(total, 1))
    from oracle import make_ref
    try:
        make_ref.load()                 # the reference's module, when oracle/_ref/ exists: imported before the fork
    except Exception:
        pass
    pool = mp.get_context("fork").Pool(procs)
    try:
        recs = [cpu_sample(n, rows, procs, pool) for _ in range(total)][args.warmup:]
    finally:
        pool.close()
    secs = sum(r["seconds"] for r in recs)
    value = procs * n * rows * len(recs) / secs
    b = dict(recs[-1], value=value)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / len(recs),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(n, args.num_steps, args.mu_per_gpu, {
                "sample": f"each bench step = one time step of {procs} mu on a {n} x {rows} band (bounded sample of the "
                          f"{args.num_steps}-step sweep, normalised per cell-update)", "cpu_count": os.cpu_count()}),
            "cpu_baseline": {k: b[k] for k in ("value", "unit", "cores", "kind", "sample", "solves_per_step")},
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
    ap.add_argument("--no-records", action="store_true", help="skip the LSPG / ECSW / training-matrix / slab sub-records")
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
    from tools import records as R
    ClockSampler = R.ClockSampler
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

    # roofline of the sweep kernel (the only kernel inside run()).  Its binding unit is the FP64 pipe: the
    # Newton sweeps of a step are fused in registers, so DRAM sees 43 B per cell-update instead of the
    # 48 K + 32 B of one HBM pass per sweep (kept below as hbm_model, SURVEY 8d's accounting).
    peaks = R.peaks()
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    abytes = algorithmic_bytes(its, cells)
    run_s = float(np.mean(run_ms)) * 1e-3
    hbm_achieved = abytes / run_s / 1e9
    dfma_peak = R.dfma_peak_tflops(dev, lib)
    dgemm_peak = R.dgemm_peak_tflops(dev)
    solves = float(its.mean()) - 1.0
    fp64_instr = NCU_FP64_INSTR_PER_CELL_UPDATE * (10.0 + 46.0 * solves + 18.0) / (10.0 + 46.0 * 4 + 18.0)   # Au/Av 10, a solve level 46, the last residual 18 (SASS counts): scaled to this run's solves
    fp64_tflops = 2e-12 * fp64_instr * (B * cells * T) / run_s
    sha = kernel_source_sha()
    stale = sha != NCU_PROFILED_SHA
    roofline = {"bound": "fp64", "achieved": fp64_tflops, "peak": dfma_peak, "unit": "TFLOP/s", "frac": fp64_tflops / dfma_peak,
                "traffic": NCU_DRAM_BYTES_PER_CELL_UPDATE * cells * T * B / max(sweeps, 1),
                "kernel": "fom_chunk_kernel" if args.kernel == "chunk" else "fom_tile_kernel",
                "peak_source": "bfd_probe_dfma in this run (DFMA issue rate, 8 chains x 32 warps per SM, best of 6)",
                "achieved_source": f"{NCU_FP64_INSTR_PER_CELL_UPDATE} FP64 warp-instruction lanes per cell-update at 4 solves (ncu "
                                   "smsp__inst_executed_pipe_fp64, profiles/r02_fom_chunk_ncu.md), scaled to this run's "
                                   f"{solves:.3f} solves per step, x 2 flop, x cell-updates / device time of run()",
                "traffic_source": f"{NCU_DRAM_BYTES_PER_CELL_UPDATE} B of DRAM traffic per cell-update (same capture), per launch",
                "ncu_constants": {"profiled_sha": NCU_PROFILED_SHA, "current_sha": sha, "stale": stale},
                "launches_per_step": sweeps, "avg_launch_us": 1e6 * run_s / max(sweeps, 1), "steps_redone": redone,
                "dgemm_peak_tflops": dgemm_peak,
                "hbm_model": {"bound": "hbm", "achieved": hbm_achieved, "peak": hbm_peak, "unit": "GB/s",
                              "frac": hbm_achieved / hbm_peak, "algorithmic_bytes_per_step": abytes,
                              "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6.65 TB/s",
                              "note": "(48*solves + 32) B per cell-update, one HBM pass per Newton sweep (SURVEY 8d); exceeds 1 "
                                      "because the fused kernel does not make those passes"}}

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
    e2e_reps = max(1, min(args.steps, 3))
    t0 = time.perf_counter()
    copy_stats = []
    for _ in range(e2e_reps):
        seen.clear()
        sweep.run(w0_np, mus_e, sink=sink)
        copy_stats.append(sweep.copy_stats or {"GBps_while_copying": 0.0, "seconds_in_copies": 0.0, "seconds_first_to_last": 0.0})
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
           "link": {"GBps_while_copying": float(np.mean([c["GBps_while_copying"] for c in copy_stats])),
                    "seconds_in_copies_per_step": float(np.mean([c["seconds_in_copies"] for c in copy_stats])),
                    "seconds_first_to_last_copy_per_step": float(np.mean([c["seconds_first_to_last"] for c in copy_stats])),
                    "worker_seconds_per_step": {k[7:]: float(np.mean([c[k] for c in copy_stats])) for k in copy_stats[0]
                                                if k.startswith("worker_")}},
           "api": f"SnapshotSweep(group={args.e2e_group}, workers={args.e2e_workers}).run(w0, mus, sink): every "
                  "(2N, T+1) trajectory copied to pinned host memory and handed to the sink"}
    del sweep

    # ---- configs [3] and [4]: timed sub-records (tools/records.py), each with its own clocks ----------------
    torch.cuda.empty_cache()
    records = {}
    if not args.no_records:
        P = R.prepare(dev) if rank == 0 else None
        if rank == 0:
            records["setup"] = {k: P[k] for k in ("fom_ms", "pod_ms", "n_snapshots", "train_steps")}
            records["lspg"] = R.lspg_record(P, dev, lib, gpu_index=local_rank)
            records["training_matrix"], records["nnls"], weights = R.training_and_weights(P, dev, lib, gpu_index=local_rank)
            records["ecsw"] = R.ecsw_record(P, dev, lib, weights, gpu_index=local_rank)
        if world > 1:
            records["lspg_sharded"] = R.lspg_sharded_record(P, dev, rank, world, dist, gpu_index=local_rank)
        P = None
        torch.cuda.empty_cache()
        records["slab"] = R.slab_record(dev, rank, world, dist, gpu_index=local_rank)
        # the same grid over 60 steps: the first ~20 steps of a trajectory from w0 = 1 need up to 10 Newton solves and
        # pay for mis-speculated counts; steps 21-60 show the steady rate
        records["slab_T60"] = R.slab_record(dev, rank, world, dist, T=60, check=0, gpu_index=local_rank)

    if rank == 0:
        cb = None
        if not args.no_cpu_baseline and world == 1:       # the CPU arm is timed at N = 1 only
            procs = host_cores()
            cb = cpu_sample(n, cpu_rows(n, procs, 20.0), procs)
            cb = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample", "solves_per_step")}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": workload_config(n, T, B, {
                    "chunk_steps": args.chunk_steps or 50, "strict": not args.non_strict, "kernel": args.kernel,
                    "l2": f"working set {B * per_mu / 2**30:.1f} GiB of snapshots per GPU, far larger than L2; no flush needed",
                    "newton_its_mean": float(its.mean()), "max_relnorm": float(rel.max())}),
                "mu_snapshots_per_s": world * B * T * args.steps / (total_ms * 1e-3),
                "mu_trajectories_per_s": world * B * args.steps / (total_ms * 1e-3),
                "roofline": roofline, "cpu_baseline": cb, "e2e": e2e, "gpu_launches": int(launches),
                "clocks": clocks.summary(), "wall_s_timed": wall, "records": records}
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
