"""Secondary measurements: LSPG PROM, ECSW HPROM and the ECSW training matrix at the
fine grid (750 x 750, k = 95 POD modes) on one B200, each with its algorithmic bytes / flops and,
with --cpu, the oracle timed on the host for a bounded sample.  Prints one JSON object per line.

The POD basis comes from FOM snapshots computed here on the device (no files): the left singular
vectors of the snapshot matrix via the Gram matrix of its columns.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(fn, reps=1):
    import torch
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    e1.synchronize()
    return out, e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=750)
    ap.add_argument("--modes", type=int, default=95)
    ap.add_argument("--train-steps", type=int, default=120)
    ap.add_argument("--lspg-steps", type=int, default=10)
    ap.add_argument("--ecsw-steps", type=int, default=100)
    ap.add_argument("--ecsw-cells", type=int, default=3000)
    ap.add_argument("--c-snaps", type=int, default=10)
    ap.add_argument("--cpu", action="store_true", help="time the oracle on the host for a bounded sample")
    a = ap.parse_args()

    import torch
    from finitedifference_b200 import _lib
    from finitedifference_b200.discretisation import make_2D_grid
    from finitedifference_b200.fom import FomSolver
    from finitedifference_b200.rom import EcswSolver, LspgSolver, ecsw_training_matrix

    dev = torch.device("cuda", 0)
    lib = _lib.get_lib()
    n, k, dt = a.grid, a.modes, 0.05
    N = n * n
    gx, gy = make_2D_grid(0, 100, 0, 100, n, n)
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    hbm = float(peaks.get("hbm_gbs", 6650.0))

    # training snapshots and POD basis on the device
    train_mu, test_mu = [[4.25, 0.015], [5.5, 0.03]], [4.75, 0.02]
    fom = FomSolver(gx, gy, dt, a.train_steps, train_mu, device=dev)
    fom.set_initial(torch.ones(2 * N, dtype=torch.float64, device=dev))
    _, fom_ms = timed(fom.run)
    S = torch.cat([fom.snapshots_device(b, layout=0) for b in range(len(train_mu))], dim=0)       # (n_snap, 2N)
    (evals, evecs), pod_ms = timed(lambda: torch.linalg.eigh(S @ S.T))
    order = torch.argsort(evals, descending=True)[:k]
    V = (S.T @ (evecs[:, order] / torch.sqrt(evals[order]))).contiguous()               # (2N, k)
    orth = float((V.T @ V - torch.eye(k, dtype=torch.float64, device=dev)).abs().max())
    w0 = np.ones(2 * N)
    print(json.dumps({"stage": "setup", "grid": n, "modes": k, "train_snapshots": int(S.shape[0]),
                      "fom_ms": fom_ms, "pod_ms": pod_ms, "basis_orthogonality_defect": orth}), flush=True)
    del fom

    # ---- LSPG PROM -----------------------------------------------------------------------
    lspg = LspgSolver(gx, gy, dt, test_mu, V, device=dev, jacobian="reference")
    lspg.run(w0, 2)                                                                       # warm-up
    lspg.res_time = lspg.jac_time = lspg.ls_time = 0.0
    l0 = lib.bfd_launch_count()
    t0 = time.perf_counter()
    (snaps, red, hist), ms = timed(lambda: lspg.run(w0, a.lspg_steps))
    wall = time.perf_counter() - t0
    its = sum(len(h) for h in hist)
    solves = sum(len(h) - 1 for h in hist)
    # per kernel, isolated
    w = snaps[a.lspg_steps].contiguous()
    wp = snaps[a.lspg_steps - 1].contiguous()
    _, jv_ms = timed(lambda: lspg.jv(lspg.jac_mesh, w, V), 5)
    _, res_ms = timed(lambda: lspg.residual(lspg.res_mesh, w, wp), 5)
    from finitedifference_b200 import rom as rom_mod
    from finitedifference_b200.rom import MeshTables
    variants = {}
    for mode, mesh in (("reference", lspg.jac_mesh), ("exact", lspg.res_mesh)):
        _, variants[f"{mode}_gather_ms"] = timed(lambda: lspg.jv(mesh, w, V, gather=True), 5)
        for rows in (1, 2, 4):
            rom_mod.JV_ROWS_PER_WARP = rows
            _, variants[f"{mode}_grid_rows{rows}_ms"] = timed(lambda: lspg.jv(mesh, w, V), 5)
        rom_mod.JV_ROWS_PER_WARP = 0
    JV = lspg.jv(lspg.jac_mesh, w, V)
    _, gram_ms = timed(lambda: JV.T @ JV, 5)
    fres = lspg.residual(lspg.res_mesh, w, wp)[0]
    _, dmma_ms = timed(lambda: lspg.normal_equations(JV, fres), 5)
    Gd, bd = lspg.normal_equations(JV, fres)
    dmma_diff = float((Gd - JV.T @ JV).abs().max() / (JV.T @ JV).abs().max())
    _, solve_ms = timed(lambda: lspg.gn_step(Gd, bd), 20)
    rom_mod.SPD_SOLVE_KERNEL = False
    _, solve_lib_ms = timed(lambda: lspg.gn_step(Gd, bd), 20)
    solve_diff = float((lspg.gn_step(Gd, bd) - (rom_mod.__setattr__("SPD_SOLVE_KERNEL", True) or lspg.gn_step(Gd, bd))).abs().max())
    _, vy_ms = timed(lambda: V @ red[1], 5)
    _, vyk_ms = timed(lambda: lspg.basis_times(V, red[1]), 5)
    vy_diff = float((lspg.basis_times(V, red[1]) - V @ red[1]).abs().max())
    jv_bytes = 8 * 2 * N * k * 2 + N * (5 * 4 + 4 * 8 + 16)        # read V once, write JV, tables + state
    print(json.dumps({
        "stage": "lspg", "steps": a.lspg_steps, "gn_residual_evals": its, "gn_solves": solves,
        "ms_per_step": ms / a.lspg_steps, "steps_per_s": 1e3 * a.lspg_steps / ms, "wall_s": wall,
        "kernels_launched": lib.bfd_launch_count() - l0,
        "rom_jv_kernel": {"ms": jv_ms, "algorithmic_bytes": jv_bytes, "GBps": jv_bytes / jv_ms / 1e6,
                          "frac_of_hbm_peak": jv_bytes / jv_ms / 1e6 / hbm, "variants": variants},
        "rom_residual_kernel": {"ms": res_ms, "algorithmic_bytes": N * (32 + 16 + 3 * 4 + 6 * 8),
                                "GBps": N * (32 + 16 + 3 * 4 + 6 * 8) / res_ms / 1e6},
        "gram_cublas": {"ms": gram_ms, "flop": 2.0 * 2 * N * k * k, "TFLOPs": 2.0 * 2 * N * k * k / gram_ms / 1e9},
        "gram_dmma_kernel": {"ms": dmma_ms, "useful_flop": 2.0 * 2 * N * k * (k + 1), "TFLOPs_useful": 2.0 * 2 * N * k * (k + 1)
                             / dmma_ms / 1e9, "blocks_computed_of_full": 90.0 / 156.0, "max_rel_diff_vs_gemm": dmma_diff,
                             "includes": "b = (JV)^T f"},
        "spd_solve_kernel": {"ms": solve_ms, "cusolver_potrf_potrs_ms": solve_lib_ms, "max_abs_diff": solve_diff},
        "V_times_y_cublas": {"ms": vy_ms, "GBps": 8 * 2 * N * k / vy_ms / 1e6},
        "basis_times_y_kernel": {"ms": vyk_ms, "GBps": 8 * 2 * N * k / vyk_ms / 1e6, "frac_of_hbm_peak": 8 * 2 * N * k / vyk_ms / 1e6 / hbm,
                                 "max_abs_diff_vs_gemv": vy_diff},
        "host_timers_s": {"jac": lspg.jac_time, "res": lspg.res_time, "ls": lspg.ls_time}}), flush=True)

    if a.cpu:
        from oracle import burgers_oracle as orc
        Vh = V.cpu().numpy()
        t0 = time.perf_counter()
        _, times, _, h1 = orc.lspg_trajectory(gx, gy, w0, dt, 1, test_mu, Vh, return_info=True)
        cpu_s = time.perf_counter() - t0
        print(json.dumps({"stage": "lspg_cpu_oracle", "steps": 1, "seconds": cpu_s, "gn_residual_evals": len(h1[0]),
                          "cores": os.cpu_count(), "gpu_gn_residual_evals_step0": len(hist[0]),
                          "first_step_rel_diff": float(np.linalg.norm(
                              orc.lspg_trajectory(gx, gy, w0, dt, 0, test_mu, Vh)[0][:, 0]
                              - snaps[0].cpu().numpy()) / np.sqrt(2 * N))}), flush=True)
    del lspg, JV, snaps

    # ---- ECSW HPROM ----------------------------------------------------------------------
    rng = np.random.default_rng(0)
    wts = np.zeros((n, n))
    inner = rng.choice((n - 2) * (n - 2), size=a.ecsw_cells, replace=False)
    ri, ci = np.divmod(inner, n - 2)
    wts[ri + 1, ci + 1] = rng.uniform(1.0, 400.0, a.ecsw_cells)
    wts[0, :] = wts[-1, :] = wts[:, 0] = wts[:, -1] = 10.0          # run_HPROM_ecsw_joshua.py bc_w
    ecsw = EcswSolver(gx, gy, wts.ravel(), dt, test_mu, V, device=dev)
    ecsw.run(w0, 2)
    l0, g0 = lib.bfd_launch_count(), ecsw.graph_kernel_launches
    t0 = time.perf_counter()
    (red_e, hist_e), ms = timed(lambda: ecsw.run(w0, a.ecsw_steps))
    wall = time.perf_counter() - t0
    print(json.dumps({"stage": "ecsw", "steps": a.ecsw_steps, "sampled_cells": int(ecsw.sample_inds.size),
                      "augmented_cells": int(ecsw.augmented_sample.size),
                      "gn_residual_evals": sum(len(h) for h in hist_e), "ms_per_step": ms / a.ecsw_steps,
                      "steps_per_s": 1e3 * a.ecsw_steps / ms, "wall_s": wall,
                      "kernels_launched": lib.bfd_launch_count() - l0 + ecsw.graph_kernel_launches - g0,
                      "note": "one CUDA-graph launch + one host look at ||f|| per Gauss-Newton iteration"}), flush=True)
    if a.cpu:
        from oracle import burgers_oracle as orc
        t0 = time.perf_counter()
        ref_e, _, _, _, h_e = orc.ecsw_trajectory(gx, gy, wts.ravel(), w0, dt, 5, test_mu, V.cpu().numpy(),
                                                  return_info=True)
        cpu_s = time.perf_counter() - t0
        diff = float(np.abs(ref_e[:, :6] - red_e[:6].T.cpu().numpy()).max() / np.abs(ref_e[:, :6]).max())
        print(json.dumps({"stage": "ecsw_cpu_oracle", "steps": 5, "seconds": cpu_s, "ms_per_step": 200 * cpu_s,
                          "max_rel_diff_reduced_state": diff}), flush=True)
    del ecsw
    from finitedifference_b200.rom import EcswBatch
    batch_mus = [[m1, m2] for m1 in (4.3, 4.75, 5.19, 5.45) for m2 in (0.017, 0.021, 0.025, 0.029)]
    t0 = time.perf_counter()
    batch = EcswBatch(gx, gy, wts.ravel(), dt, batch_mus, V, device=dev)
    setup_s = time.perf_counter() - t0
    batch.run(w0, 2)                                                                      # warm-up
    t0 = time.perf_counter()
    out = batch.run(w0, a.ecsw_steps)
    wall = time.perf_counter() - t0
    print(json.dumps({"stage": "ecsw_batch", "parameters": len(batch_mus), "steps_each": a.ecsw_steps,
                      "wall_s": wall, "hprom_steps_per_s": len(batch_mus) * a.ecsw_steps / wall,
                      "setup_and_capture_s": setup_s,
                      "gn_residual_evals": sum(len(h) for _, hist in out for h in hist)}), flush=True)
    del batch

    # ---- ECSW training matrix ------------------------------------------------------------
    fom = FomSolver(gx, gy, dt, a.c_snaps + 1, [test_mu], device=dev)
    fom.set_initial(torch.ones(2 * N, dtype=torch.float64, device=dev))
    fom.run()
    traj = fom.snapshots_device(0, layout=0)                                                       # (c_snaps+2, 2N)
    cur, prev = traj[1:a.c_snaps + 1].T.contiguous(), traj[0:a.c_snaps].T.contiguous()
    ecsw_training_matrix(cur[:, :1], prev[:, :1], V, gx, gy, dt, test_mu, device=dev, out_device=True)
    l0 = lib.bfd_launch_count()
    Cm, ms = timed(lambda: ecsw_training_matrix(cur, prev, V, gx, gy, dt, test_mu, device=dev, out_device=True))
    c_bytes = a.c_snaps * (8 * k * N + 8 * 2 * N * k)               # per snapshot: V read once, the C block written (fused kernel)
    print(json.dumps({"stage": "ecsw_training_matrix", "snapshots": a.c_snaps, "shape": list(Cm.shape),
                      "ms": ms, "ms_per_snapshot": ms / a.c_snaps, "bytes_model": c_bytes,
                      "GBps": c_bytes / ms / 1e6, "frac_of_hbm_peak": c_bytes / ms / 1e6 / hbm,
                      "kernels_launched": lib.bfd_launch_count() - l0}), flush=True)
    if a.cpu:
        from oracle import burgers_oracle as orc
        t0 = time.perf_counter()
        Co = orc.ecsw_training_matrix(cur[:, :1].cpu().numpy(), prev[:, :1].cpu().numpy(), V.cpu().numpy(), gx, gy,
                                      dt, test_mu)
        cpu_s = time.perf_counter() - t0
        print(json.dumps({"stage": "ecsw_training_matrix_cpu_oracle", "snapshots": 1, "seconds": cpu_s,
                          "max_abs_diff": float(np.abs(Co - Cm[:k].cpu().numpy()).max())}), flush=True)


if __name__ == "__main__":
    main()
