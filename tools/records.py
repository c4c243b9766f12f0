"""Timed sub-records of bench.py: BASELINE.json configs [3] (LSPG PROM, ECSW HPROM, training matrix at
the fine grid with k = 95) and [4] (one large grid over several GPUs by slabs), plus the row-sharded
LSPG.  Every record is timed with CUDA events on the launching stream and carries the clocks sampled
while it ran.  Importable (bench.py) and runnable:

    python tools/records.py                       # one GPU: lspg, training_matrix, nnls, ecsw
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \\
        --master-port 29533 tools/records.py      # + slab and lspg_sharded

Inputs are produced by the pipeline itself, the way run_prom.py / run_HPROM_ecsw_joshua.py do at full
size (tools/run_config4.py) but bounded so that the default bench still ends in minutes: a POD basis
(k = 95, 'svd') of two training trajectories of `train_steps` steps, the ECSW training matrix of
`c_snaps` snapshots of the training parameter, weights by the 20-block NNLS with the interior mask
and bc_w = 10.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

DT = 0.05
TEST_MU = [4.75, 0.02]
TRAIN_MUS = [[4.25, 0.0225], [5.5, 0.03]]      # the ECSW training parameter (run_HPROM_ecsw_joshua.py:43) first


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        return {}


def timed(fn, reps=1):
    """(result, ms per call) with CUDA events on the current stream."""
    import torch
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    e1.synchronize()
    return out, e0.elapsed_time(e1) / reps


def dfma_peak_tflops(dev, lib):
    """DFMA issue rate of this device, now (bfd_probe_dfma): best of 5 launches of ~4 ms."""
    import torch
    blocks, iters = 148 * 2, 1 << 14
    sink = torch.empty(blocks * 1024, dtype=torch.float64, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    best = 0.0
    for _ in range(6):
        _, ms = timed(lambda: lib.bfd_probe_dfma(sink.data_ptr(), blocks, iters, st))
        best = max(best, 2.0 * 8 * iters * blocks * 1024 / (ms * 1e-3) / 1e12)
    return best


def dgemm_peak_tflops(dev, n=4096):
    """cuBLAS Dgemm n^3 on this device (the FP64 tensor-pipe denominator SURVEY 8d asks for)."""
    import torch
    a = torch.randn((n, n), dtype=torch.float64, device=dev)
    b = torch.randn((n, n), dtype=torch.float64, device=dev)
    c = torch.empty((n, n), dtype=torch.float64, device=dev)
    torch.mm(a, b, out=c)
    best = 0.0
    for _ in range(4):
        _, ms = timed(lambda: torch.mm(a, b, out=c))
        best = max(best, 2.0 * n ** 3 / (ms * 1e-3) / 1e12)
    return best


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while a record is being timed."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index, period_ms=100):
        self.index, self.rows, self.proc, self.period = index, [], None, period_ms

    def __enter__(self):
        import subprocess
        import threading
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", str(self.period)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
            t0 = time.perf_counter()                 # nvidia-smi needs a moment: the record starts once it samples
            while not self.rows and time.perf_counter() - t0 < 3.0:
                time.sleep(0.01)
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *exc):
        if self.proc:
            time.sleep(1.5e-3 * self.period)         # one more sample after the work
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


def prepare(dev, n=750, k=95, train_steps=120):
    """Grid, training trajectories (device, time-major) and the POD basis of config [3], all on `dev`."""
    import torch
    from finitedifference_b200.discretisation import make_2D_grid
    from finitedifference_b200.fom import FomSolver
    from finitedifference_b200.pod import pod_tensor
    gx, gy = make_2D_grid(0, 100, 0, 100, n, n)
    N = n * n
    fom = FomSolver(gx, gy, DT, train_steps, TRAIN_MUS, device=dev)
    fom.set_initial(torch.ones(2 * N, dtype=torch.float64, device=dev))
    _, fom_ms = timed(fom.run)
    traj = [fom.snapshots_device(b, layout=0) for b in range(len(TRAIN_MUS))]          # (T+1, 2N) each
    S = torch.cat(traj, dim=0).T.contiguous()                                           # (2N, n_snap) as POD() takes it
    (u, sig), pod_ms = timed(lambda: pod_tensor(S, k, "svd"))
    V = u[:, :k].contiguous()
    del S, u, fom
    return {"gx": gx, "gy": gy, "n": n, "k": k, "V": V, "train": traj[0], "fom_ms": fom_ms, "pod_ms": pod_ms,
            "n_snapshots": sum(t.shape[0] for t in traj), "train_steps": train_steps}


def lspg_record(P, dev, lib, steps=10, gpu_index=0):
    """LSPG PROM (inviscid_burgers_implicit2D_LSPG, :133-200) at the test parameter: ms per time step
    and its kernels timed alone, each against the roof that bounds it."""
    import torch
    from finitedifference_b200.rom import LspgSolver
    n, k, V = P["n"], P["k"], P["V"]
    N = n * n
    hbm = float(peaks().get("hbm_gbs", 6650.0))
    w0 = np.ones(2 * N)
    lspg = LspgSolver(P["gx"], P["gy"], DT, TEST_MU, V, device=dev, jacobian="reference")
    lspg.run(w0, 2)
    with ClockSampler(gpu_index, 50) as clk:
        (snaps, red, hist), ms = timed(lambda: lspg.run(w0, steps))
        w, wp = snaps[steps].contiguous(), snaps[steps - 1].contiguous()
        JV, jv_ms = timed(lambda: lspg.jv(lspg.jac_mesh, w, V), 10)
        (f, _), res_ms = timed(lambda: lspg.residual(lspg.res_mesh, w, wp), 10)
        (G, b), gram_ms = timed(lambda: lspg.normal_equations(JV, f), 10)
        _, solve_ms = timed(lambda: lspg.gn_step(G, b), 20)
        _, vy_ms = timed(lambda: lspg.basis_times(V, red[1]), 10)
        dgemm = dgemm_peak_tflops(dev)
    jv_bytes = 8 * 2 * N * k * 2                      # V read once, J V written
    gram_exec = 2.0 * (2 * N) * 64 * 90               # DMMA flop executed: 90 blocks of 8 x 8 per row of J V (78 of G on or above the diagonal + 12 of b)
    rec = {"steps": steps, "ms_per_step": ms / steps, "gn_residual_evals": sum(len(h) for h in hist),
           "gn_solves": sum(len(h) - 1 for h in hist), "jacobian": "reference (:167 operator)",
           "kernels": {
               "rom_jv_grid_kernel": {"ms": jv_ms, "algorithmic_bytes": jv_bytes, "GBps": jv_bytes / jv_ms / 1e6,
                                      "frac_of_hbm": jv_bytes / jv_ms / 1e6 / hbm},
               "gram_partial+reduce (bfd_gram)": {"ms": gram_ms, "dmma_tflops_executed": gram_exec / gram_ms / 1e9,
                                                  "useful_tflops": 2.0 * 2 * N * k * (k + 1) / gram_ms / 1e9,
                                                  "dgemm_peak_tflops": dgemm,
                                                  "frac_of_dgemm_peak": gram_exec / gram_ms / 1e9 / dgemm},
               "basis_times_y (V y)": {"ms": vy_ms, "GBps": 8 * 2 * N * k / vy_ms / 1e6,
                                       "frac_of_hbm": 8 * 2 * N * k / vy_ms / 1e6 / hbm},
               "rom_residual_kernel": {"ms": res_ms}, "spd_solve_kernel": {"ms": solve_ms}},
           "clocks": clk.summary()}
    del lspg, JV, snaps
    return rec


def training_and_weights(P, dev, lib, c_snaps=10, blocks=20, bc_w=10.0, gpu_index=0):
    """compute_ECSW_training_matrix_2D (:2795-2816) on `c_snaps` snapshots of the training parameter and
    the weights of run_HPROM_ecsw_joshua.py:66-99 (interior mask, 20 NNLS blocks, bc_w on the boundary).
    Returns (training-matrix record, nnls record, weights)."""
    import torch
    from finitedifference_b200.nnls import ecsw_weights
    from finitedifference_b200.rom import ecsw_training_matrix
    n, k, V, S = P["n"], P["k"], P["V"], P["train"]
    N = n * n
    hbm = float(peaks().get("hbm_gbs", 6650.0))
    T = S.shape[0] - 1
    stride = max(1, (T - 3) // c_snaps)
    cur = S[3:T + 1:stride][:c_snaps].T.contiguous()              # snaps[:, 3::stride], prev = snaps[:, 0::stride] (:58-59)
    prev = S[0:T - 2:stride][:c_snaps].T.contiguous()
    ns = cur.shape[1]
    ecsw_training_matrix(cur[:, :1], prev[:, :1], V, P["gx"], P["gy"], DT, TRAIN_MUS[0], device=dev, out_device=True)
    with ClockSampler(gpu_index, 50) as clk:
        Cm, ms = timed(lambda: ecsw_training_matrix(cur, prev, V, P["gx"], P["gy"], DT, TRAIN_MUS[0], device=dev,
                                                    out_device=True))
    c_bytes = ns * (8 * k * N + 8 * 2 * N * k)                     # per snapshot: the C block written, V read once
    rec_c = {"snapshots": ns, "shape": list(Cm.shape), "ms_per_snapshot": ms / ns, "algorithmic_bytes": c_bytes,
             "GBps": c_bytes / ms / 1e6, "frac_of_hbm": c_bytes / ms / 1e6 / hbm, "kernel": "bfd_ecsw_c_grid",
             "clocks": clk.summary()}
    idxs = np.zeros((n, n), dtype=bool)
    idxs[1:-1, 1:-1] = True
    interior = torch.from_numpy(np.flatnonzero(idxs.ravel())).to(dev)
    t0 = time.perf_counter()
    Ci = Cm[:, interior].contiguous()
    wts = ecsw_weights(Ci, blocks)
    nnls_s = time.perf_counter() - t0
    wt = torch.from_numpy(wts).to(dev)
    tgt = Ci.sum(dim=1)
    resid = float(torch.linalg.norm(Ci @ wt - tgt) / torch.linalg.norm(tgt))     # the figure :88-90 prints
    full = np.full((n, n), bc_w)
    full[idxs] = wts
    rec_n = {"rows": int(Ci.shape[0]), "columns": int(Ci.shape[1]), "blocks": blocks, "seconds": nnls_s,
             "nnls_solver_residual": resid, "N_e": int(np.sum(full > 0)), "interior_cells_selected": int(np.sum(wts > 0)),
             "algorithm": "Lawson-Hanson on the device (finitedifference_b200/nnls.py)"}
    del Cm, Ci
    return rec_c, rec_n, full.ravel()


def ecsw_record(P, dev, lib, weights, steps=100, gpu_index=0):
    """ECSW HPROM (inviscid_burgers_ecsw_fixed, :202-273) on pipeline-produced weights: ms per time step."""
    from finitedifference_b200.rom import EcswSolver
    n = P["n"]
    w0 = np.ones(2 * n * n)
    ecsw = EcswSolver(P["gx"], P["gy"], weights, DT, TEST_MU, P["V"], device=dev)
    ecsw.run(w0, 2)
    l0, g0 = lib.bfd_launch_count(), ecsw.graph_kernel_launches
    with ClockSampler(gpu_index, 50) as clk:
        (red, hist), ms = timed(lambda: ecsw.run(w0, steps))
    evals = sum(len(h) for h in hist)
    return {"steps": steps, "ms_per_step": ms / steps, "steps_per_s": 1e3 * steps / ms,
            "sampled_cells": int(ecsw.sample_inds.size), "augmented_cells": int(ecsw.augmented_sample.size),
            "gn_residual_evals": evals, "us_per_gn_iteration": 1e3 * ms / max(evals, 1),
            "kernels_launched": int(lib.bfd_launch_count() - l0 + ecsw.graph_kernel_launches - g0),
            "bound": "latency (one CUDA-graph launch and one host look at the norm per Gauss-Newton iteration)",
            "clocks": clk.summary()}


def slab_record(dev, rank, world, dist, grid=8192, T=20, dx=100.0 / 750.0, check=512, gpu_index=0, kernel="chunk"):
    """BASELINE.json config [4]: one `grid`^2 FOM trajectory over `world` GPUs by slab decomposition
    (peer stores over NVLink inside the sweep kernel, finitedifference_b200/slab.py).  Strong scaling;
    time = max over ranks of the device time of run().  With `check`, a smaller grid is first compared
    bit for bit with a single-GPU solve.  The cell size is the fine grid's: at dx = 100/8192, dt = 0.05
    the reference's own Newton iteration diverges."""
    import torch
    from finitedifference_b200.discretisation import make_2D_grid
    from finitedifference_b200.fom import FomSolver
    from finitedifference_b200.slab import FomSlab, connect_distributed
    barrier = (lambda: dist.barrier()) if world > 1 else (lambda: None)

    def make(n, steps):
        gx, gy = make_2D_grid(0, n * dx, 0, n * dx, n, n)
        s = FomSlab(gx, gy, DT, steps, [TEST_MU], rank, world, device=dev, kernel=kernel)
        connect_distributed(s) if world > 1 else s.connect()
        return gx, gy, s

    def trajectory(s):
        s.init()
        barrier()
        s.set_initial(torch.ones(2 * s.ny * s.nx_local, dtype=torch.float64, device=dev))
        barrier()
        _, ms = timed(s.run)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    exact = None
    if check:
        gx, gy, s = make(check, 6)
        trajectory(s)
        mine = s.snapshots_device(0)
        solo = FomSolver(gx, gy, DT, 6, [TEST_MU], device=dev, kernel=kernel)
        solo.set_initial(torch.ones(2 * check * check, dtype=torch.float64, device=dev))
        solo.run()
        whole = solo.snapshots_device(0, layout=0).reshape(7, 2, check, check)[:, :, :, s.c0:s.c1]
        ok = torch.tensor([int(bool(torch.equal(mine, whole)))], device=dev)
        if world > 1:
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        exact = bool(ok.item())
        s.close()
        del solo, mine, whole
        torch.cuda.empty_cache()
        barrier()
    gx, gy, s = make(grid, T)
    trajectory(s)                                                   # warm-up
    with ClockSampler(gpu_index, 100) as clk:
        ms = float(np.mean([trajectory(s) for _ in range(2)]))
    its, rel = s.diagnostics()
    rec = {"grid": grid, "num_steps": T, "n_gpus": world, "scaling": "strong", "kernel": kernel, "ms_per_trajectory": ms,
           "cell_updates_per_s": grid * grid * T / (ms * 1e-3), "newton_its_mean": float(its.mean()),
           "max_relnorm": float(rel.max()), "steps_redone": int(s.steps_redone), "dx": dx,
           "bit_exact_vs_single_gpu": exact, "check_grid": check,
           "halo_bytes_per_trajectory": int((world - 1) * T * grid * (2 * float(its.mean()) + 1) * 16),
           "clocks": clk.summary()}
    s.close()
    torch.cuda.empty_cache()
    barrier()
    return rec


def lspg_sharded_record(P_rank0, dev, rank, world, dist, n=750, k=95, steps=10, gpu_index=0):
    """LSPG PROM with the grid rows dealt to the ranks and one all-reduce of the k x (k+1) normal equations
    per Gauss-Newton iteration.  Rank 0 broadcasts its basis."""
    import torch
    from finitedifference_b200.discretisation import make_2D_grid
    from finitedifference_b200.rom import ShardedLspgSolver
    gx, gy = make_2D_grid(0, 100, 0, 100, n, n)
    N = n * n
    V = P_rank0["V"] if rank == 0 else torch.empty((2 * N, k), dtype=torch.float64, device=dev)
    if world > 1:
        dist.broadcast(V, 0)
    Vh = V.cpu().numpy()
    if rank != 0:
        del V
    w0 = np.ones(2 * N)
    solver = ShardedLspgSolver(gx, gy, DT, TEST_MU, Vh, device=dev, jacobian="reference")
    solver.run(w0, 2)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    with ClockSampler(gpu_index, 50) as clk:
        (snaps, red, hist), ms = timed(lambda: solver.run(w0, steps))
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    rec = {"n_gpus": world, "steps": steps, "ms_per_step": float(t.item()) / steps,
           "gn_residual_evals": sum(len(h) for h in hist), "rows_per_rank": int(solver.row_range[1] - solver.row_range[0]),
           "basis_fraction_held": float(solver.state_cells.size) / N, "red_checksum": float(red.abs().sum().item()),
           "collective": "all-reduce of [G | b] (k x (k+1) doubles) and of ||f||^2 per Gauss-Newton iteration (NCCL)",
           "clocks": clk.summary()}
    del solver
    torch.cuda.empty_cache()
    return rec


def main():
    import torch
    import torch.distributed as dist
    from finitedifference_b200 import _lib
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.get_lib()
    out = {"n_gpus": world}
    P = prepare(dev) if rank == 0 else None
    if rank == 0:
        out["setup"] = {k: P[k] for k in ("fom_ms", "pod_ms", "n_snapshots", "train_steps")}
        out["dfma_peak_tflops"] = dfma_peak_tflops(dev, lib)
        out["lspg"] = lspg_record(P, dev, lib, gpu_index=local)
        out["training_matrix"], out["nnls"], weights = training_and_weights(P, dev, lib, gpu_index=local)
        out["ecsw"] = ecsw_record(P, dev, lib, weights, gpu_index=local)
    if world > 1:
        out["lspg_sharded"] = lspg_sharded_record(P, dev, rank, world, dist, gpu_index=local)
    P = None
    torch.cuda.empty_cache()
    out["slab"] = slab_record(dev, rank, world, dist, gpu_index=local)
    out["slab_T60"] = slab_record(dev, rank, world, dist, T=60, check=0, gpu_index=local)
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
