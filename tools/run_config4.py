"""BASELINE.json config [3] end to end on one B200, at the fine grid, with the inputs the reference's own
scripts produce -- nothing leaves HBM between the stages:

  run_prom.py:59-81           9-mu training sweep (3 x 3 lattice) + the test parameter  -> FomSolver (one batch)
                              POD(snapshots, 95, 'rsvd', random_state=0) on the 1 125 000 x 4 509 matrix
  run_HPROM_ecsw_joshua.py    training matrix on [:, 3:500:10] / [:, 0:497:10] of mu = (4.25, 0.0225) (4750 x 562 500,
    :55-100                   21 GB), interior mask, 20-block NNLS, bc_w = 10
  run_prom.py:97-110          LSPG PROM, 500 steps at mu = (4.75, 0.02); error against the FOM
  run_HPROM_ecsw_joshua.py    ECSW HPROM, 500 steps; error against the FOM
    :104-127

Prints one JSON object with the seconds of every stage (CUDA events / wall clock as marked), the clocks
sampled over the whole run and the figures the reference's scripts print (NNLS residual :88-90, N_e :101,
relative errors).  Sizes can be reduced for a quick run (--grid, --steps, --snap-factor).
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=750)
    ap.add_argument("--steps", type=int, default=500)
    ap.add_argument("--modes", type=int, default=95)
    ap.add_argument("--snap-factor", type=int, default=10)
    ap.add_argument("--blocks", type=int, default=20)
    ap.add_argument("--pod", default="rsvd", choices=["rsvd", "svd"])
    ap.add_argument("--lspg-steps", type=int, default=0, help="0 = all")
    ap.add_argument("--selector", default="nnls", choices=["nnls", "ecm"])
    a = ap.parse_args()

    import torch
    from finitedifference_b200 import hypernet2D as h
    from finitedifference_b200.ecm import ecm_weights
    from finitedifference_b200.fom import FomSolver
    from finitedifference_b200.nnls import ecsw_weights
    from finitedifference_b200.pod import pod_tensor
    from finitedifference_b200.rom import EcswSolver, LspgSolver, ecsw_training_matrix
    from tools.records import ClockSampler, timed

    dev = torch.device("cuda", 0)
    n, T, k, dt = a.grid, a.steps, a.modes, 0.05
    N = n * n
    gx, gy = h.make_2D_grid(0, 100, 0, 100, n, n)
    mu1 = np.linspace(4.25, 5.5, 3)
    mu2 = np.linspace(0.015, 0.03, 3)
    train = [[float(m1), float(m2)] for m1 in mu1 for m2 in mu2]            # config.py:25-27, run_prom.py:59-71
    test_mu, ecsw_mu = [4.75, 0.02], [4.25, 0.0225]
    out = {"grid": n, "steps": T, "modes": k, "stages": {}}
    st = out["stages"]
    with ClockSampler(0, 200) as clk:
        wall0 = time.perf_counter()
        fom = FomSolver(gx, gy, dt, T, train + [test_mu], device=dev)
        fom.set_initial(torch.ones(2 * N, dtype=torch.float64, device=dev))
        _, ms = timed(fom.run)
        st["fom_sweep_10_mu"] = {"seconds": ms * 1e-3, "cell_updates_per_s": 10 * N * T / (ms * 1e-3)}
        S, ms = timed(lambda: torch.cat([fom.snapshots_device(m, layout=1) for m in range(len(train))], dim=1))
        st["snapshot_matrix"] = {"seconds": ms * 1e-3, "shape": list(S.shape), "GB": S.numel() * 8 / 1e9}
        ib = train.index(ecsw_mu)
        Sb = fom.snapshots_device(ib, layout=1)                                       # (2N, T+1) of the ECSW training parameter
        hdm = fom.snapshots_device(len(train), layout=1)                             # the test parameter's FOM
        del fom
        torch.cuda.empty_cache()
        (u, sig), ms = timed(lambda: pod_tensor(S, k, a.pod, random_state=0))
        V = u[:, :k].contiguous()
        st["pod"] = {"seconds": ms * 1e-3, "method": a.pod, "sigma_k_over_sigma_1": float(sig[k - 1] / sig[0]),
                     "orthogonality_defect": float((V.T @ V - torch.eye(k, dtype=torch.float64, device=dev)).abs().max())}
        del S, u
        torch.cuda.empty_cache()

        # LSPG PROM (run_prom.py:97-110)
        w0 = np.ones(2 * N)
        Ls = a.lspg_steps or T
        lspg = LspgSolver(gx, gy, dt, test_mu, V, device=dev, jacobian="reference")
        (snaps, red, hist), ms = timed(lambda: lspg.run(w0, Ls))
        err = (torch.linalg.norm(snaps.T - hdm[:, :Ls + 1], dim=0) / torch.linalg.norm(snaps.T, dim=0))     # compute_error :3150
        st["lspg"] = {"seconds": ms * 1e-3, "steps": Ls, "ms_per_step": ms / Ls, "gn_residual_evals": sum(len(x) for x in hist),
                      "relative_error_mean": float(err.mean()), "relative_error_max": float(err.max())}
        del lspg, snaps
        torch.cuda.empty_cache()

        # training matrix and weights (run_HPROM_ecsw_joshua.py:55-100)
        f = a.snap_factor
        cur, prev = Sb[:, 3:T:f].contiguous(), Sb[:, 0:T - 3:f].contiguous()
        Cm, ms = timed(lambda: ecsw_training_matrix(cur, prev, V, gx, gy, dt, ecsw_mu, device=dev, out_device=True))
        st["training_matrix"] = {"seconds": ms * 1e-3, "shape": list(Cm.shape), "GB": Cm.numel() * 8 / 1e9,
                                 "ms_per_snapshot": ms / cur.shape[1]}
        idxs = np.zeros((n, n), dtype=bool)
        idxs[1:-1, 1:-1] = True
        Ci = Cm[:, torch.from_numpy(np.flatnonzero(idxs.ravel())).to(dev)].contiguous()
        del Cm
        torch.cuda.empty_cache()
        t0 = time.perf_counter()
        wts = ecsw_weights(Ci, a.blocks) if a.selector == "nnls" else ecm_weights(Ci)[0]
        sel_s = time.perf_counter() - t0
        tgt = Ci.sum(dim=1)
        resid = float(torch.linalg.norm(Ci @ torch.from_numpy(wts).to(dev) - tgt) / torch.linalg.norm(tgt))
        full = np.full((n, n), 10.0)
        full[idxs] = wts
        st["weights"] = {"seconds_wall": sel_s, "selector": a.selector, "blocks": a.blocks, "solver_residual": resid,
                         "N_e": int(np.sum(full > 0)), "interior_cells_selected": int(np.sum(wts > 0))}
        del Ci
        torch.cuda.empty_cache()

        # ECSW HPROM (run_HPROM_ecsw_joshua.py:104-127)
        ecsw = EcswSolver(gx, gy, full.ravel(), dt, test_mu, V, device=dev)
        ecsw.run(w0, 2)
        (redh, hh), ms = timed(lambda: ecsw.run(w0, T))
        rom = V @ redh.T                                                             # basis_trunc @ rom_y (:127)
        err = torch.linalg.norm(rom - hdm, dim=0) / torch.linalg.norm(rom, dim=0)
        st["ecsw_hprom"] = {"seconds": ms * 1e-3, "steps": T, "ms_per_step": ms / T, "gn_residual_evals": sum(len(x) for x in hh),
                            "sampled_cells": int(ecsw.sample_inds.size), "relative_error_mean": float(err.mean()),
                            "relative_error_max": float(err.max())}
        out["wall_seconds_total"] = time.perf_counter() - wall0
    out["clocks"] = clk.summary()
    out["reference_log_figures"] = "run_prom.py POD 650-890 s; FOM 2.44e4 s per mu; LSPG ~30 s/step (SURVEY.md 3.2, output_54914041.log)"
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
