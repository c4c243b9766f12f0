/*
 * burgers_b200.h -- C ABI of the B200-native (sm_100a) implicit 2D inviscid-Burgers hot path.
 *
 * The reference (SADPR/FiniteDifference) has no FFI layer: its boundary is the module-level
 * Python functions of BurgersFD_Clean{Coarse,Fine,TestAE}/hypernet2D.py.  The Python shim
 * finitedifference_b200/hypernet2D.py keeps those names and signatures and calls the entry
 * points below through ctypes.  Every entry point names the reference function it replaces.
 *
 * Conventions
 *   - all pointers are raw DEVICE pointers unless the name starts with `h_` (host);
 *   - all real data is float64, index sets are int64;
 *   - "natural" state layout is the reference's: w = [u.ravel(), v.ravel()], each (ny, nx)
 *     row-major, cell id p = r*nx + c (run_fom.py:33-35, hypernet2D.py:2614);
 *   - nothing allocates: callers size a workspace with the matching *_workspace_bytes();
 *   - every function returns 0 on success, a negative BFD_E* code otherwise, and never throws;
 *   - `stream` is a cudaStream_t passed as void*.
 */
#ifndef BURGERS_B200_H
#define BURGERS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BFD_OK 0
#define BFD_EINVAL (-1)       /* bad argument */
#define BFD_ECUDA (-2)        /* CUDA runtime error, see bfd_last_cuda_error() */
#define BFD_EWORKSPACE (-3)   /* workspace too small */
#define BFD_ETIMEOUT (-4)     /* a wavefront hand-off timed out on the device */
#define BFD_ENOTCONVERGED (-5)
#define BFD_ETOOBIG (-6)      /* size beyond a fixed-capacity entry point's stated maximum, see bfd_last_cuda_error() */

/* Library / device identification. */
int bfd_abi_version(void);
const char *bfd_last_cuda_error(void);

/* ------------------------------------------------------------------------------------------
 * Full-order model: replaces inviscid_burgers_implicit2D (hypernet2D.py:72-131) together with
 * newton_raphson (:1887-1933), inviscid_burgers_res2D_alt (:2588-2646),
 * inviscid_burgers_exact_jac2D (:2703-2732) and scipy's spsolve, batched over `batch`
 * parameter samples mu.
 *
 * Discretisation inputs (device, float64), computed by the caller exactly as the reference does:
 *   cx[nx]   = 0.5*dt*(1/dx[c])           diagonal of 0.5*dt*Dx        (make_ddx, :2486-2492)
 *   cxl[nx]  = 0.5*dt*(1/dx[c-1]), 0 at c=0   minus the sub-diagonal
 *   cy[ny], cyd[ny]                        same along y
 *   src[batch][nx] = dt*0.02*exp(mu2*xc)   source term             (:2625)
 *   lbc[batch][ny] = 0.5*dt*mu1^2/dx       inlet ghost flux, applied at c == 0 (:2628-2631)
 * ------------------------------------------------------------------------------------------ */
#define BFD_FOM_KMAX 6         /* Newton solves fused into one pass over a time step */
#define BFD_FOM_MAX_CHUNK 64   /* time steps advanced by one launch */

typedef struct bfd_fom_desc {
    int32_t nx, ny;          /* cells in x (fast axis) and y */
    int32_t batch;           /* number of mu samples advanced together */
    int32_t num_steps;       /* T; T+1 snapshots are kept per mu */
    int32_t max_its;         /* Newton cap; the reference passes 100 (:123) */
    int32_t chunk_steps;     /* time steps per launch, 1..BFD_FOM_MAX_CHUNK; 0 = default (50) */
    int32_t newton_hint;     /* first guess of the Newton solves per step, 1..BFD_FOM_KMAX; 0 = 4 */
    int32_t strict;          /* 1: iteration counts and iterates exactly as newton_raphson decides
                                them (a step whose speculated solve count was too high is redone);
                                0: such a step keeps its further-converged iterate */
    int32_t uniform_y;       /* 1: cy[r] and cyd[r>0] are one value (held in a register) */
    int32_t kernel;          /* 0 (default): fom_chunk_kernel, the Newton levels of a step split over the warps of a
                                CTA; 1: fom_tile_kernel, one warp per strip and time step with all levels in registers,
                                three time steps of a strip in flight per CTA (no CTA barrier, no data-movement warp;
                                measured 34.5-35.2 G against the default's 39.4 G cell-updates/s on B200, kept selectable: DESIGN.md 4.1b).  Same results. */
    double rel_tol;          /* relnorm_cutoff, < 1; the reference passes 1e-12 (:123) */
} bfd_fom_desc;

/* Device pointers to the discretisation inputs listed above. */
typedef struct bfd_fom_coef {
    const double *cx, *cxl, *cy, *cyd, *src, *lbc;
} bfd_fom_coef;

size_t bfd_fom_workspace_bytes(const bfd_fom_desc *d);

/* Lay the workspace out and reset the per-sample control blocks (asynchronous on `stream`). */
int bfd_fom_init(const bfd_fom_desc *d, void *workspace, size_t workspace_bytes, void *stream);

/* Initial condition, natural layout.  w0_stride = 0 broadcasts one (2N) state to every mu,
 * otherwise w0 + m*w0_stride is sample m's state. */
int bfd_fom_set_initial(const bfd_fom_desc *d, void *workspace, const double *w0,
                        size_t w0_stride, void *stream);

/* Advance every sample num_steps steps (Newton to rel_tol each step); returns when done.
 * h_pinned: >= 64 bytes of page-locked host memory used to poll completion.  When the memory is mapped into
 * the device's address space (cudaHostAlloc / cudaHostRegister memory is: cudaHostGetDevicePointer succeeds) the
 * GPU writes its status words there itself after every launch (one-thread kernel, no copy engine: a small
 * device-to-host copy would wait behind whatever bulk copies the process has in flight); otherwise they are copied.
 * On return h_pinned[0..3] (int32) = {samples finished, device error code, kernels launched,
 * time steps that were redone or continued because the speculated solve count was wrong}. */
int bfd_fom_run(const bfd_fom_desc *d, void *workspace, const bfd_fom_coef *coef, void *h_pinned,
                void *stream);

/* Copy out snapshots t0 <= t < t1 of sample `mu_index`.
 * layout 0: out[(t-t0)*2N + dof]         (time-major, one contiguous state per snapshot)
 * layout 1: out[dof*ld + (t-t0)]         (the reference's (2N, T+1) C-order array, ld >= t1-t0) */
int bfd_fom_export(const bfd_fom_desc *d, const void *workspace, int mu_index, int t0, int t1,
                   int layout, size_t ld, double *out, void *stream);

/* Per-sample, per-step diagnostics: number of residual-norm evaluations (== len(resnorms) of
 * newton_raphson) and the last relative residual.  its: int32 [batch][num_steps];
 * relnorm: float64 [batch][num_steps]. Device pointers. */
int bfd_fom_diagnostics(const bfd_fom_desc *d, const void *workspace, int32_t *its,
                        double *relnorm, void *stream);

/* Pipeline trace of the most recent launches: for every (sample, time step of the launch,
 * strip) the %globaltimer (ns) at which the strip started and finished that time step.
 * out: uint64 [batch][chunk_steps][nstrips][3], device pointer (third word: hand-off blocks that
 * had not arrived when first looked at << 32 | spin iterations); nstrips = ceil(nx/32).  Entries
 * of steps a launch did not run keep older values.  Profiling aid (see profiles/). */
int bfd_fom_trace(const bfd_fom_desc *d, const void *workspace, uint64_t *out, void *stream);

/* Number of kernels launched by this library since load (for bench.py's gpu_launches). */
uint64_t bfd_launch_count(void);

/* ------------------------------------------------------------------------------------------
 * Natural-layout operators shared by the PROM / HPROM paths and the tests.
 * ------------------------------------------------------------------------------------------ */

/* r = residual(w; wp) -- inviscid_burgers_res2D / _res2D_alt (:2544-2646).
 * Also accumulates sum(r^2) into *norm2 (device, must be zeroed by the caller) when non-NULL. */
int bfd_residual(int nx, int ny, const double *cx, const double *cxl, const double *cy,
                 const double *cyd, const double *src, const double *lbc, const double *w,
                 const double *wp, double *r, double *norm2, void *stream);

/* DFMA issue-rate probe (measurement aid for bench.py's FP64 roofline; no counterpart in the reference):
 * `blocks` CTAs of 1024 threads run 8 independent chains of `iters` DFMAs each and write one double per
 * thread to sink[blocks * 1024].  flop = 2 * 8 * iters * blocks * 1024; the caller times the launch. */
int bfd_probe_dfma(double *sink, int blocks, int iters, void *stream);

/* ------------------------------------------------------------------------------------------
 * One grid over several GPUs: slab decomposition along x (second row).
 *
 * The grid's 32-column strips are dealt to the ranks in contiguous blocks; each rank (one process
 * per GPU) holds only its strips' snapshots.  Upwinding makes the coupling one-directional: the
 * last strip of rank g hands its right-edge records (the same self-validating records strips on
 * one GPU exchange) straight into the halo buffer of rank g+1 with peer stores over NVLink,
 * tile by tile inside the running kernel -- there is no separate exchange step and no
 * collective.  The per-sample control blocks, residual norms and verdicts of a slab run live in
 * rank 0's workspace ("home"); every rank updates them with system-scope atomics.
 *
 * `d` describes the GLOBAL problem.  Arrays passed to the slab entry points are LOCAL to the
 * slab: coef->cx, cxl [nx_local], coef->src [batch][nx_local], w0 / exported snapshots in the
 * natural layout of a (ny x nx_local) grid; coef->cy, cyd, lbc are per row and global.
 * `home` and `next_halo` are addresses valid in THIS process: cudaIpc mappings of the other
 * ranks' workspaces (bfd_ipc_*), or plain pointers when the slabs live in one process.
 * Call order on every rank: _init; (all ranks: stream sync + barrier); _set_initial; _run.
 * ------------------------------------------------------------------------------------------ */
typedef struct bfd_fom_slab {
    int32_t rank, world;             /* this slab and the number of slabs */
    int32_t strip_begin, strip_end;  /* strips [begin, end) of the global grid held here */
    void *home;                      /* rank 0's workspace (NULL or own workspace on rank 0) */
    void *next_halo;                 /* rank+1's workspace + bfd_fom_slab_halo_offset(rank+1); NULL on the last rank */
} bfd_fom_slab;

size_t bfd_fom_slab_workspace_bytes(const bfd_fom_desc *d, const bfd_fom_slab *slab);
size_t bfd_fom_slab_halo_offset(const bfd_fom_desc *d, const bfd_fom_slab *slab);
int bfd_fom_slab_init(const bfd_fom_desc *d, const bfd_fom_slab *slab, void *workspace, size_t workspace_bytes,
                      void *stream);
int bfd_fom_slab_set_initial(const bfd_fom_desc *d, const bfd_fom_slab *slab, void *workspace, const double *w0,
                             size_t w0_stride, void *stream);
int bfd_fom_slab_run(const bfd_fom_desc *d, const bfd_fom_slab *slab, void *workspace, const bfd_fom_coef *coef,
                     void *h_pinned, void *stream);
int bfd_fom_slab_export(const bfd_fom_desc *d, const bfd_fom_slab *slab, const void *workspace, int mu_index,
                        int t0, int t1, int layout, size_t ld, double *out, void *stream);
int bfd_fom_slab_diagnostics(const bfd_fom_desc *d, const bfd_fom_slab *slab, const void *workspace, int32_t *its,
                             double *relnorm, void *stream);

/* Device memory that can be mapped into another process on the same node (cudaMalloc +
 * cudaIpcGetMemHandle / OpenMemHandle; peer access is enabled by the mapping). */
#define BFD_IPC_HANDLE_BYTES 64
int bfd_mem_alloc(size_t bytes, void **ptr);
int bfd_mem_free(void *ptr);
int bfd_ipc_export(void *ptr, unsigned char handle[BFD_IPC_HANDLE_BYTES]);
int bfd_ipc_open(const unsigned char handle[BFD_IPC_HANDLE_BYTES], void **ptr);
int bfd_ipc_close(void *ptr);

/* ------------------------------------------------------------------------------------------
 * Reduced-order models (LSPG PROM on the whole grid, ECSW/ECM HPROM on the sampled mesh).
 *
 * A "mesh" is a list of rows (cells whose residual is evaluated) with, per row, the state cells
 * its stencil reads.  State arrays hold `ncells` cells per component: u of state cell q at
 * w[q], v at w[ncells + q] (the whole grid, or the augmented sample [aug | N+aug] of
 * inviscid_burgers_ecsw_fixed, hypernet2D.py:242-245).  Index tables are int32, -1 = no such
 * neighbour (contributes zero, as the first row of make_ddx does, :2491):
 *   p          own cell
 *   xl, yd     left (r, c-1) and lower-row (r-1, c) cells: residual (:2544-2586, :2648-2701)
 *   ja, jb     the two cells of the JACOBIAN's y-difference: (p, yd) for the exact operator
 *              (get_ops, :2509-2520); the transposed cells (c, r), (c, r-1) for the operator
 *              inviscid_burgers_implicit2D_LSPG builds at :167 (rows-only permutation)
 *   cx, cxl, cy, cyd   per-row entries of 0.5*dt*Dx / 0.5*dt*Dy (diagonal, minus sub-diagonal)
 *   src, lbc   per-row source and inlet terms (:2625-2631, :2660-2669)
 *   weight     per-row ECSW weight (gauss_newton_ECSW_2D, :2012) or NULL
 * All pointers are device pointers.
 * ------------------------------------------------------------------------------------------ */
typedef struct bfd_rom_mesh {
    int64_t nrows, ncells;
    const int32_t *p, *xl, *yd, *ja, *jb;
    const double *cx, *cxl, *cy, *cyd, *src, *lbc, *weight;
} bfd_rom_mesh;

/* r[0:nrows] = u-residual rows, r[nrows:2*nrows] = v-residual rows, times weight when given;
 * adds sum(r^2) to *norm2 (device, zeroed by the caller) when non-NULL.
 * Replaces inviscid_burgers_res2D (:2544-2586) / _res2D_ecsw (:2648-2701). */
int bfd_rom_residual(const bfd_rom_mesh *mesh, const double *w, const double *wp, double *r,
                     double *norm2, void *stream);

/* JV[2*nrows][k] (row-major) = rows of J(w) @ V, times weight when given; V is the basis
 * restricted to the state cells, row-major [2*ncells][k].  Replaces inviscid_burgers_exact_jac2D
 * (:2703-2732) / _exact_jac2D_ecsw (:2734-2744) followed by J.dot(basis) (:1993, :2030, :2812). */
int bfd_rom_jv(const bfd_rom_mesh *mesh, const double *w, const double *V, int k, double *JV,
               void *stream);

/* The same J V for a mesh that is the WHOLE nx x ny grid in row order (nrows == ncells == nx*ny,
 * p[i] == i, no weights), computed by walking the grid instead of gathering: left neighbours stay
 * in registers and a warp that owns `rows_per_warp` (1, 2 or 4; 0 = default) grid rows shares the
 * y-difference cells between them, so the basis is read ~1.25x (exact operator) instead of 4x.
 * transposed = 0: the exact operator of get_ops (:2509-2520), ja = p, jb = yd;
 * transposed = 1: the operator inviscid_burgers_implicit2D_LSPG builds at :167 (nx == ny),
 *                 ja = (c, r), jb = (c, r-1).
 * Only the coefficient tables cx, cxl, cy, cyd of `mesh` are read.  Same arithmetic, in the same
 * order, as bfd_rom_jv.  Replaces inviscid_burgers_exact_jac2D (:2703-2732) + J.dot(basis) (:1993). */
int bfd_rom_jv_grid(const bfd_rom_mesh *mesh, int nx, int ny, int transposed, int rows_per_warp,
                    const double *w, const double *V, int k, double *JV, void *stream);

/* bfd_rom_jv_grid on a SLAB of grid rows [r_lo, r_hi) (LSPG rows dealt to the ranks of a process
 * group): `mesh` lists the owned cells in row order (nrows = nx*(r_hi-r_lo)); the state and basis
 * come as two slabs that start one grid row below the owned rows, row0 = max(r_lo-1, 0):
 *   wP, VP : grid rows [row0, r_hi) in row order, [u | v] with nx*(r_hi-row0) cells per component;
 *   wA, VA : (transposed = 1 only) the transposed cells as a column slab: cell (c, r), r in [row0, r_hi),
 *            at c*(r_hi-row0) + (r-row0), [u | v] with nx*(r_hi-row0) cells per component.
 * JV[2*nrows][k]: rows of the owned cells [u | v].  Same arithmetic as bfd_rom_jv_grid. */
int bfd_rom_jv_grid_slab(const bfd_rom_mesh *mesh, int nx, int ny, int transposed, int r_lo, int r_hi,
                         const double *wP, const double *VP, const double *wA, const double *VA, int k,
                         double *JV, void *stream);

/* Values of the Jacobian rows of `mesh` at state w, 12 per row i at vals[12*i ..], in fixed slots
 * whose columns the caller derives from the index tables (N = ncells):
 *   u row i:      (p) (xl) (ja) (jb) (N+ja) (N+jb)
 *   v row nrows+i:(p) (xl) (N+p) (N+xl) (N+ja) (N+jb)
 * Slots of an absent neighbour hold 0.  Replaces the sparse products and bmat of
 * inviscid_burgers_exact_jac2D (:2703-2732) / _exact_jac2D_ecsw (:2734-2744); weight is ignored. */
int bfd_jac_csr_values(const bfd_rom_mesh *mesh, const double *w, double *vals, void *stream);

/* One snapshot's block of the ECSW training matrix (compute_ECSW_training_matrix_2D, :2807-2814):
 * C[j*ldc + cell] = r[cell]*JV[cell][j] + r[nrows+cell]*JV[nrows+cell][j], j < k, cell < nrows. */
int bfd_ecsw_c_block(int64_t nrows, int k, const double *r, const double *JV, double *C, size_t ldc,
                     void *stream);

/* The same block straight from the basis: C[j*ldc + cell] = r[cell]*(J(w)V)[cell][j] +
 * r[nrows+cell]*(J(w)V)[nrows+cell][j] without materialising J V (r from bfd_rom_residual on an
 * unweighted mesh).  Replaces one snapshot's iteration of compute_ECSW_training_matrix_2D
 * (:2805-2814: residual, exact Jacobian, J.dot(basis), the cell loop). */
int bfd_ecsw_c_fused(const bfd_rom_mesh *mesh, const double *w, const double *r, const double *V, int k,
                     double *C, size_t ldc, void *stream);

/* bfd_ecsw_c_fused for a mesh that is the whole nx x ny grid in row order (see bfd_rom_jv_grid):
 * the grid is walked in segments of 8 cells, left neighbours in registers, the 8 x k tile of
 *   C[j*ldc + cell] = r[cell]*(J(w)V)[cell][j] + r[nrows+cell]*(J(w)V)[nrows+cell][j]
 * transposed through shared memory and written along cells.  Exact operator of get_ops only (what
 * compute_ECSW_training_matrix_2D is called with, run_HPROM_ecsw_joshua.py:61-64).
 * Replaces one snapshot's iteration of compute_ECSW_training_matrix_2D (:2805-2814). */
int bfd_ecsw_c_grid(const bfd_rom_mesh *mesh, int nx, int ny, const double *w, const double *r,
                    const double *V, int k, double *C, size_t ldc, void *stream);

/* Normal equations of a Gauss-Newton step on the FP64 tensor cores (DMMA):
 *   G[k][k] = X^T X (row-major, both triangles),  b[k] = X^T f,   X row-major [M][k], k <= 96.
 * Replaces numpy.linalg.lstsq(JV, -f) of gauss_newton_LSPG (:1994) / gauss_newton_ECSW_2D (:2038)
 * together with a k x k Cholesky solve of G dy = -b by the caller.  Only the blocks on or above
 * the diagonal are computed; partial sums are added in a fixed order (deterministic).
 * workspace: bfd_gram_workspace_bytes(M, k) bytes of device memory (0 = unsupported k). */
size_t bfd_gram_workspace_bytes(int64_t M, int k);
int bfd_gram(const double *X, const double *f, int64_t M, int k, double *G, double *b, void *workspace,
             size_t workspace_bytes, void *stream);

/* bfd_gram without a workspace argument and without any allocation, for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * This entry point has NO counterpart in the reference.
 * The partial sums go to a fixed device array inside the library
 * (BFD_GRAM_STATIC_PARTS x 96 x 104 doubles = 11.3 MiB per GPU that loads the library, held until
 * the process exits), one CTA per SM.  Limits: M <= BFD_GRAM_STATIC_MAX_ROWS, k <= 96; anything
 * larger returns BFD_ETOOBIG with the reason in bfd_last_cuda_error().  Calls no cudaMalloc /
 * cudaFree / allocating library routine.  All calls share the one array: calls that overlap on
 * the GPU (two streams without an ordering between them) race -- serialise them.  Nothing needs
 * clearing between calls (every entry read is overwritten first).  Same G and b as bfd_gram up to
 * the summation order of the partials (fewer, longer partials: agreement ~1e-15 relative). */
#define BFD_GRAM_STATIC_PARTS 148
#define BFD_GRAM_STATIC_MAX_ROWS (1 << 21)
int bfd_gram_static(const double *X, const double *f, int64_t M, int k, double *G, double *b, void *stream);

/* x = scale * G^-1 b for a symmetric positive definite k x k matrix (row-major, lower triangle
 * read), k <= 96: Cholesky + two triangular solves in one single-CTA launch (a matrix that is not
 * positive definite yields NaNs, as silent as the reference's lstsq is about rank).  With bfd_gram
 * this is the Gauss-Newton step dy = lstsq(JV, -f) of :1994 / :2038 (scale = -1). */
int bfd_spd_solve(const double *G, const double *b, int k, double scale, double *x, void *stream);

/* w[M] = V y for a row-major basis V[M][k], k <= 96 (basis.dot(y) of :1999, :2042, :196): HBM-bound
 * streaming product, four rows per warp. */
int bfd_basis_times_y(const double *V, const double *y, int64_t M, int k, double *w, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* BURGERS_B200_H */
