"""Large golden fixtures: the UNMODIFIED reference at the sizes BASELINE.json's configs name.

Runs only in the build container (needs /root/reference; see gen_golden.py for how the reference
module is imported).  Stages are separate invocations so that they can run on different cores;
full trajectories are parked under --scratch (not committed), only small samples go to
tests/golden/.

    python oracle/gen_golden_large.py fom64          # 64^2, 500 steps (~20 s)
    python oracle/gen_golden_large.py fom250a        # 250^2, 500 steps, mu = (5.19, 0.026)  (~15 min)
    python oracle/gen_golden_large.py fom250b        # 250^2, 500 steps, mu = (4.25, 0.0225) (~15 min)
    python oracle/gen_golden_large.py hat250         # LSPG, analytic basis, 250^2, k = 95
    python oracle/gen_golden_large.py hat750         # LSPG, analytic basis, 750^2, k = 95 (minutes)
    python oracle/gen_golden_large.py rom250         # needs fom250a+b: POD, LSPG, training matrix,
                                                     # NNLS weights the reference's way, ECSW HPROM

Fixtures written (float64 / int64, np.savez_compressed):
  fom_n64_T500.npz      norms of all 501 snapshots, every 25th snapshot in full, all snapshots at every
                        8th dof, Newton counts  (BurgersFD_CleanCoarse/run_fom.py:9 parameters)
  fom_n250_T500.npz     the same at 250^2 with coarser samples (config 1 of BASELINE.json)
  fom_n250_T500_train.npz   norms + counts of the ECSW training parameter's trajectory
  lspg_hat_n250.npz / lspg_hat_n750.npz   LSPG steps on oracle.fixtures.hat_basis: reduced
                        coordinates, GN counts, strided states
  rom_n250_k95.npz      POD coefficients (W/sigma, 252 x 95), sample of the basis, LSPG steps, one
                        training-matrix block (strided), NNLS weights (run_HPROM_ecsw_joshua.py:55-100),
                        index sets, ECSW HPROM reduced coordinates, GN counts
"""
import argparse
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
from oracle import fixtures as fx                     # noqa: E402
from oracle.gen_golden import OUT, import_reference, quiet, record_counts   # noqa: E402

DT = 0.05
MU_A = [5.19, 0.026]      # BurgersFD_CleanCoarse/run_fom.py:9
MU_B = [4.25, 0.0225]     # ECSW training parameter, run_HPROM_ecsw_joshua.py:43
MU_TEST = [4.75, 0.02]    # run_prom.py / run_HPROM_ecsw_joshua.py default query point


def fom(ref, n, steps, mu):
    gx, gy = ref.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    counts, restore = record_counts(ref, "newton_raphson")
    t0 = time.time()
    with quiet():
        snaps = ref.inviscid_burgers_implicit2D(gx, gy, w0, DT, steps, mu)
    restore()
    print(f"fom n={n} mu={mu}: {time.time() - t0:.1f} s", flush=True)
    return snaps, np.array(counts)


def save_fom(name, n, mu, snaps, counts, dof_stride, t_stride):
    np.savez_compressed(os.path.join(OUT, name), n=n, mu=mu, dt=DT, steps=snaps.shape[1] - 1,
                        norms=np.linalg.norm(snaps, axis=0), newton_its=counts,
                        dof_stride=dof_stride, t_stride=t_stride,
                        dof_sample=snaps[::dof_stride, :].copy(), t_sample=snaps[:, ::t_stride].copy())


def lspg_hat(ref, n, steps, name):
    gx, gy = ref.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    V = fx.hat_basis(n)
    counts, restore = record_counts(ref, "gauss_newton_LSPG")
    t0 = time.time()
    with quiet():
        snaps, times = ref.inviscid_burgers_implicit2D_LSPG(gx, gy, w0, DT, steps, MU_TEST, V)
    restore()
    print(f"lspg hat n={n}: {time.time() - t0:.1f} s, GN counts {counts}", flush=True)
    stride = 997
    np.savez_compressed(os.path.join(OUT, name), n=n, mu=MU_TEST, dt=DT, steps=steps, k=V.shape[1],
                        red_coords=V.T @ snaps, gn_its=np.array(counts), num_its=times[0],
                        stride=stride, sample=snaps[::stride, :].copy(),
                        norms=np.linalg.norm(snaps, axis=0), basis_sample=V[::stride, :].copy())


def rom250(ref, scratch):
    from scipy.optimize import nnls
    n, k, T = 250, 95, 500
    gx, gy = ref.make_2D_grid(0, 100, 0, 100, n, n)
    w0 = np.ones(2 * n * n)
    Sa = np.load(os.path.join(scratch, "fom250a.npy"))
    Sb = np.load(os.path.join(scratch, "fom250b.npy"))
    S = fx.pod_training_matrix([Sa, Sb])
    # POD(S, method='svd') (hypernet2D.py:2746-2771) also yields W and sigma; keep W/sigma
    u, s, vh = np.linalg.svd(S, full_matrices=False)
    u_ref, s_ref = ref.POD(S, method="svd")
    assert np.array_equal(s, s_ref)
    coef = np.ascontiguousarray(vh[:k].T / s[:k][None, :])
    V = fx.pod_from_coef(S, coef)
    print("basis: |S coef - U| max", np.abs(V - u_ref[:, :k]).max(), "sigma_k/sigma_1", s[k - 1] / s[0], flush=True)
    out = dict(n=n, k=k, dt=DT, mu_train=[MU_A, MU_B], mu=MU_TEST, coef=coef, sigma=s,
               basis_stride=499, basis_sample=V[::499, :].copy())

    # LSPG PROM, 5 steps (hypernet2D.py:133-200)
    counts, restore = record_counts(ref, "gauss_newton_LSPG")
    t0 = time.time()
    with quiet():
        snaps, times = ref.inviscid_burgers_implicit2D_LSPG(gx, gy, w0, DT, 5, MU_TEST, V)
    restore()
    print(f"lspg: {time.time() - t0:.1f} s, GN counts {counts}", flush=True)
    out.update(lspg_steps=5, lspg_red=V.T @ snaps, lspg_gn_its=np.array(counts), lspg_num_its=times[0],
               lspg_stride=251, lspg_sample=snaps[::251, :].copy(), lspg_norms=np.linalg.norm(snaps, axis=0))

    # training matrix the way run_HPROM_ecsw_joshua.py:55-64 builds it (training mu = MU_B)
    t0 = time.time()
    with quiet():
        C = ref.compute_ECSW_training_matrix_2D(Sb[:, 3:T:10], Sb[:, 0:T - 3:10], V, ref.inviscid_burgers_res2D,
                                                ref.inviscid_burgers_exact_jac2D, gx, gy, DT, MU_B)
    print(f"training matrix {C.shape}: {time.time() - t0:.1f} s", flush=True)
    # block of snapshot 7 (columns strided) and of snapshot 31 (a contiguous run of cells)
    out.update(C_snap=[7, 31], C_stride=53, C_block7=C[7 * k:8 * k, ::53].copy(),
               C_block31=C[31 * k:32 * k, 30000:30600].copy(), C_colsum=C.sum(axis=0), C_rowsum=C.sum(axis=1))

    # weights: interior mask, 20 blocks of scipy nnls, bc_w = 10 (run_HPROM_ecsw_joshua.py:66-99)
    idxs = np.zeros((n, n))
    idxs[1:-1, 1:-1] = 1
    Ci = np.ascontiguousarray(C[:, (idxs == 1).ravel()], dtype=np.float64)
    del C
    t0 = time.time()
    from joblib import Parallel, delayed
    res = Parallel(n_jobs=int(os.environ.get("NNLS_JOBS", "4")))(
        delayed(nnls)(c, c.sum(axis=1), maxiter=9999999999) for c in np.array_split(Ci, 20, axis=1))
    wts = np.hstack([r[0] for r in res])
    nnls_res = np.linalg.norm(Ci @ wts - Ci.sum(axis=1)) / np.linalg.norm(Ci.sum(axis=1))
    print(f"nnls: {time.time() - t0:.1f} s, residual {nnls_res:.3e}, N_e {np.sum(wts > 0)}", flush=True)
    full = 10.0 * np.ones((n, n))
    full[idxs > 0] = wts
    weights = full.ravel()
    sample_inds = np.where(weights != 0)[0]
    aug = ref.generate_augmented_mesh(gx, gy, sample_inds)

    # ECSW HPROM, 20 steps (hypernet2D.py:202-273)
    counts, restore = record_counts(ref, "gauss_newton_ECSW_2D")
    t0 = time.time()
    with quiet():
        red, _ = ref.inviscid_burgers_ecsw_fixed(gx, gy, weights, w0, DT, 20, MU_TEST, V)
    restore()
    print(f"ecsw: {time.time() - t0:.1f} s, GN counts {counts}", flush=True)
    out.update(weights=weights, nnls_residual=nnls_res, sample_inds=sample_inds.astype(np.int64),
               augmented_sample=np.asarray(aug, dtype=np.int64), ecsw_steps=20, ecsw_red=red,
               ecsw_gn_its=np.array(counts))
    np.savez_compressed(os.path.join(OUT, "rom_n250_k95.npz"), **out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("stage")
    ap.add_argument("--scratch", default="/tmp/bfd_gold")
    args = ap.parse_args()
    os.makedirs(args.scratch, exist_ok=True)
    ref = import_reference()
    if args.stage == "fom64":
        S, c = fom(ref, 64, 500, MU_A)
        save_fom("fom_n64_T500.npz", 64, MU_A, S, c, 8, 25)
    elif args.stage == "fom250a":
        S, c = fom(ref, 250, 500, MU_A)
        np.save(os.path.join(args.scratch, "fom250a.npy"), S)
        save_fom("fom_n250_T500.npz", 250, MU_A, S, c, 251, 250)
    elif args.stage == "fom250b":
        S, c = fom(ref, 250, 500, MU_B)
        np.save(os.path.join(args.scratch, "fom250b.npy"), S)
        np.savez_compressed(os.path.join(OUT, "fom_n250_T500_train.npz"), n=250, mu=MU_B, dt=DT, steps=500,
                            norms=np.linalg.norm(S, axis=0), newton_its=c)
    elif args.stage == "hat250":
        lspg_hat(ref, 250, 6, "lspg_hat_n250.npz")
    elif args.stage == "hat750":
        lspg_hat(ref, 750, 4, "lspg_hat_n750.npz")
    elif args.stage == "rom250":
        rom250(ref, args.scratch)
    else:
        raise SystemExit("unknown stage " + args.stage)


if __name__ == "__main__":
    main()
