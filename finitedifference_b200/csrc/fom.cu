// Full-order model: batched Newton sweeps as fused residual + block forward substitution.
//
// Replaces, for the FOM path of the reference (BurgersFD_CleanFine/hypernet2D.py):
//   inviscid_burgers_implicit2D :72-131, newton_raphson :1887-1933,
//   inviscid_burgers_res2D_alt :2588-2646, inviscid_burgers_exact_jac2D :2703-2732 and the
//   SuperLU spsolve at :1930.
//
// One Newton iteration J d = r, x <- x - d never builds J: with first-order upwind
// differences the Jacobian is 2x2-block lower triangular (each cell couples to itself, its
// left and its lower-row neighbour), so the solve is a block forward substitution
//     D_p d_p = r_p - A_L d_L - A_D d_D
// in which the cells of an anti-diagonal are independent.  A warp owns a strip of 32*CPL
// columns and walks that wavefront: lane l is l rows behind lane 0, the left neighbour's
// (u, v, fluxes, d) arrive by __shfl_up, the upper neighbour's are the lane's own previous
// step.  States live in the skewed strip layout (common.cuh) so the anti-diagonal a warp
// touches per step is one contiguous line of HBM.  Strips are chained left to right through
// self-describing 64-byte records in global memory (no flags, no fences: a record is valid
// when it no longer holds the sentinel; the consumer re-arms it).
//
// The sweep also accumulates ||r(x_k)||^2, so "residual norm, test, solve" of newton_raphson
// costs one pass: 48 B per cell (read x_k 16, read wp 16, write x_{k+1} 16), 32 B when
// x_k == wp (first sweep of a step).  The last strip of a sample to finish applies the
// reference's stopping rule on the device and sets up the sample's next sweep or next time
// step, so samples progress independently and the host only keeps the launch queue full.
#include "wavefront.cuh"

namespace bfd {

struct FomCtl {            // one per mu sample, 64 bytes
    int t;                 // time level being advanced (t -> t+1); == num_steps when finished
    int k;                 // Newton sweep index inside the step
    int in_slot;           // pool slot holding x_k
    int out_slot;          // pool slot receiving x_{k+1}
    int wp_slot;           // pool slot holding snapshot t
    int slot_p, slot_q;    // the two work slots of this step (x_1 -> P, x_2 -> Q, x_3 -> P, ...)
    int next_fresh;        // next never-used pool slot
    unsigned arrive;       // strips of this sample that finished the current sweep
    int pad0;
    double init_norm;      // ||r(x_0)|| of the current step
    double pad1[2];
};
static_assert(sizeof(FomCtl) == 64, "FomCtl must stay 64 bytes");

struct FomGlobal {         // one per workspace, 64 bytes
    unsigned ticket;       // virtual block id dispenser (wraps every launch)
    int done;              // samples that completed all steps
    int error;             // BFD_E* raised on the device
    int pad[13];
};

struct FomLayout {         // byte offsets inside the workspace
    Ssl ssl;
    int batch, num_steps;
    size_t off_global, off_ctl, off_snap_slot, off_its, off_relnorm, off_partial, off_edge,
        off_pool, total;
    size_t edge_stride_mu;   // doubles
    size_t pool_stride_mu;   // doubles
};

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

static FomLayout fom_layout(const bfd_fom_desc *d) {
    FomLayout L;
    L.ssl = make_ssl(d->nx, d->ny, d->cells_per_lane);
    L.batch = d->batch;
    L.num_steps = d->num_steps;
    size_t o = 0;
    L.off_global = o;    o = align_up(o + sizeof(FomGlobal), 256);
    L.off_ctl = o;       o = align_up(o + sizeof(FomCtl) * d->batch, 256);
    L.off_snap_slot = o; o = align_up(o + sizeof(int) * (size_t)d->batch * (d->num_steps + 1), 256);
    L.off_its = o;       o = align_up(o + sizeof(int) * (size_t)d->batch * d->num_steps, 256);
    L.off_relnorm = o;   o = align_up(o + sizeof(double) * (size_t)d->batch * d->num_steps, 256);
    L.off_partial = o;   o = align_up(o + sizeof(double) * (size_t)d->batch * L.ssl.nstrips, 256);
    L.edge_stride_mu = (size_t)L.ssl.nstrips * d->ny * kEdgeDoubles;
    L.off_edge = o;      o = align_up(o + sizeof(double) * L.edge_stride_mu * d->batch, 256);
    L.pool_stride_mu = L.ssl.state_doubles * (size_t)(d->num_steps + 2);
    L.off_pool = o;      o = align_up(o + sizeof(double) * L.pool_stride_mu * d->batch, 256);
    L.total = o;
    return L;
}

static int check_desc(const bfd_fom_desc *d) {
    if (!d || d->nx < 1 || d->ny < 1 || d->batch < 1 || d->num_steps < 0 || d->max_its < 1) return BFD_EINVAL;
    if (d->cells_per_lane != 1 && d->cells_per_lane != 2) return BFD_EINVAL;
    return BFD_OK;
}

// Arguments of the sweep kernel (passed by value).
struct FomArgs {
    Ssl ssl;
    int batch, num_steps, max_its;
    double rel_tol;
    const double *cx, *cxl, *cy, *cyd, *src, *lbc;
    FomGlobal *glob;
    FomCtl *ctl;
    int *snap_slot;     // [batch][num_steps+1]
    int *its;           // [batch][num_steps]
    double *relnorm;    // [batch][num_steps]
    double *partial;    // [batch][nstrips]
    double *edge;       // [batch][nstrips][ny][kEdgeDoubles]
    double *pool;       // [batch][num_steps+2][state]
    size_t edge_stride_mu, pool_stride_mu;
};

static FomArgs fom_args(const bfd_fom_desc *d, const FomLayout &L, void *ws) {
    char *b = (char *)ws;
    FomArgs a;
    a.ssl = L.ssl;
    a.batch = d->batch; a.num_steps = d->num_steps; a.max_its = d->max_its; a.rel_tol = d->rel_tol;
    a.cx = a.cxl = a.cy = a.cyd = a.src = a.lbc = nullptr;
    a.glob = (FomGlobal *)(b + L.off_global);
    a.ctl = (FomCtl *)(b + L.off_ctl);
    a.snap_slot = (int *)(b + L.off_snap_slot);
    a.its = (int *)(b + L.off_its);
    a.relnorm = (double *)(b + L.off_relnorm);
    a.partial = (double *)(b + L.off_partial);
    a.edge = (double *)(b + L.off_edge);
    a.pool = (double *)(b + L.off_pool);
    a.edge_stride_mu = L.edge_stride_mu;
    a.pool_stride_mu = L.pool_stride_mu;
    return a;
}

}  // namespace bfd

namespace bfd {

// One Newton sweep of every still-active sample.  grid = ceil(batch*nstrips / WARPS) CTAs of
// WARPS warps; warp <-> (sample m, strip j).  Virtual block ids come from a ticket so that a
// strip's left neighbour (a smaller id) is always already resident when the strip waits on it.
template <int CPL, int PF, int WARPS>
__global__ void __launch_bounds__(WARPS * 32) fom_sweep_kernel(const FomArgs a) {
    BFD_SHARED(unsigned, s_vb);
    BFD_SHARED_ARRAY(double, s_edge, [WARPS][2][kEdgeSmemDoubles]);
    if (threadIdx.x == 0) s_vb = atomicInc(&a.glob->ticket, gridDim.x - 1);
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nstrips = a.ssl.nstrips;
    const int item = (int)s_vb * WARPS + warp;
    if (item >= a.batch * nstrips) return;
    const int m = item / nstrips, j = item - m * nstrips;
    FomCtl *ctl = a.ctl + m;
    const int t = ctl->t;
    if (t >= a.num_steps) return;  // this sample has finished its trajectory

    constexpr int W = 32 * CPL;
    const int nx = a.ssl.nx, ny = a.ssl.ny;
    const int in_slot = ctl->in_slot, wp_slot = ctl->wp_slot, out_slot = ctl->out_slot;
    const bool first = (in_slot == wp_slot);  // x_0 == wp: read the state once
    const size_t sd = a.ssl.state_doubles;
    const size_t strip_off = (size_t)j * a.ssl.strip_stride + (size_t)lane * CPL;
    const double *pool_m = a.pool + (size_t)m * a.pool_stride_mu;
    const double *xin = pool_m + (size_t)in_slot * sd + strip_off;
    const double *xwp = pool_m + (size_t)wp_slot * sd + strip_off;
    double *xout = a.pool + (size_t)m * a.pool_stride_mu + (size_t)out_slot * sd + strip_off;
    const int last_row = a.ssl.slot_rows - 1;
    const int nsteps = ny + kSkew;

    // per-lane column constants
    double cx[CPL], cxl[CPL], srcq[CPL];
    bool colok[CPL];
#pragma unroll
    for (int q = 0; q < CPL; ++q) {
        const int c = j * W + lane * CPL + q;
        colok[q] = c < nx;
        const int cc = colok[q] ? c : nx - 1;
        cx[q] = a.cx[cc];
        cxl[q] = a.cxl[cc];
        srcq[q] = a.src[(size_t)m * nx + cc];
    }
    const bool inlet = (j == 0 && lane == 0);
    const double *lbc = a.lbc + (size_t)m * ny;

    // hand-off records: read from the strip on the left, written for the strip on the right
    double *edge_m = a.edge + (size_t)m * a.edge_stride_mu;
    EdgeReader ein;
    EdgeWriter eout;
    ein.init(edge_m + (size_t)(j > 0 ? j - 1 : 0) * ny * kEdgeDoubles, s_edge[warp][0], ny, lane, j > 0);
    eout.init(edge_m + (size_t)j * ny * kEdgeDoubles, s_edge[warp][1], ny, lane, j < nstrips - 1);

    auto load = [&](int s, double (&u)[CPL], double (&v)[CPL], double (&up)[CPL], double (&vp)[CPL]) {
        const size_t o = (size_t)(s < last_row ? s : last_row) * (2 * W);
        if constexpr (CPL == 1) {
            u[0] = __ldg(xin + o);
            v[0] = __ldg(xin + o + W);
            if (!first) { up[0] = __ldg(xwp + o); vp[0] = __ldg(xwp + o + W); }
        } else {
            const double2 a0 = __ldg(reinterpret_cast<const double2 *>(xin + o));
            const double2 a1 = __ldg(reinterpret_cast<const double2 *>(xin + o + W));
            u[0] = a0.x; u[1] = a0.y; v[0] = a1.x; v[1] = a1.y;
            if (!first) {
                const double2 b0 = __ldg(reinterpret_cast<const double2 *>(xwp + o));
                const double2 b1 = __ldg(reinterpret_cast<const double2 *>(xwp + o + W));
                up[0] = b0.x; up[1] = b0.y; vp[0] = b1.x; vp[1] = b1.y;
            }
        }
    };

    // prefetch ring (registers): rows s .. s+PF-1
    double bu[PF][CPL], bv[PF][CPL], bup[PF][CPL], bvp[PF][CPL];
#pragma unroll
    for (int i = 0; i < PF; ++i) load(i, bu[i], bv[i], bup[i], bvp[i]);

    // own cells of the previous step = upper neighbours (r-1, c) of this step
    Nbr prev[CPL];
#pragma unroll
    for (int q = 0; q < CPL; ++q) prev[q] = nbr_zero();
    double nrm = 0.0;

    for (int s0 = 0; s0 < nsteps; s0 += PF) {
#pragma unroll
        for (int i = 0; i < PF; ++i) {
            const int s = s0 + i;
            if (s >= nsteps) break;  // warp-uniform
            double u[CPL], v[CPL], up[CPL], vp[CPL];
#pragma unroll
            for (int q = 0; q < CPL; ++q) {
                u[q] = bu[i][q]; v[q] = bv[i][q];
                up[q] = first ? u[q] : bup[i][q];
                vp[q] = first ? v[q] : bvp[i][q];
            }
            load(s + PF, bu[i], bv[i], bup[i], bvp[i]);

            const int err = ein.step(s);
            if (err && lane == 0) atomicExch(&a.glob->error, err);

            // left neighbour (r, c-1): lane-1's last cell of the previous step
            Nbr L;
            L.u = shfl_up1(prev[CPL - 1].u); L.v = shfl_up1(prev[CPL - 1].v);
            L.F = shfl_up1(prev[CPL - 1].F); L.G = shfl_up1(prev[CPL - 1].G);
            L.du = shfl_up1(prev[CPL - 1].du); L.dv = shfl_up1(prev[CPL - 1].dv);
            L.H = 0.0;
            if (lane == 0) {
                L = nbr_zero();
                if (j > 0 && s < ny) ein.get(s, L);
            }

            const int r = s - lane;
            const bool rowok = (unsigned)r < (unsigned)ny;
            const int rr = rowok ? r : 0;
            const double cy = __ldg(a.cy + rr), cyd = __ldg(a.cyd + rr);
            const double lb = (inlet && rowok) ? __ldg(lbc + rr) : 0.0;

            double ou[CPL], ov[CPL];
#pragma unroll
            for (int q = 0; q < CPL; ++q) {
                double r2;
                Nbr c = cell_update(u[q], v[q], up[q], vp[q], cx[q], cxl[q], cy, cyd, srcq[q],
                                    q == 0 ? lb : 0.0, L, prev[q], r2);
                if (rowok && colok[q]) nrm += r2;
                else c = nbr_zero();          // outside the grid: contributes nothing, stays zero
                ou[q] = c.u - c.du;
                ov[q] = c.v - c.dv;
                prev[q] = c;
                L = c;                        // left neighbour of column q+1
            }

            double *o = xout + (size_t)s * (2 * W);
            if constexpr (CPL == 1) {
                o[0] = ou[0];
                o[W] = ov[0];
            } else {
                *reinterpret_cast<double2 *>(o) = make_double2(ou[0], ou[1]);
                *reinterpret_cast<double2 *>(o + W) = make_double2(ov[0], ov[1]);
            }
            eout.step(s - 31, prev[CPL - 1]);
        }
    }

    // ||r||^2 of this strip, fixed reduction order (deterministic iteration counts)
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) nrm += __shfl_xor_sync(0xffffffffu, nrm, off);
    if (lane != 0) return;
    a.partial[(size_t)m * nstrips + j] = nrm;
    __threadfence();
    if (atomicInc(&ctl->arrive, (unsigned)nstrips - 1) != (unsigned)nstrips - 1) return;
    __threadfence();

    // Last strip of this sample: newton_raphson's bookkeeping (hypernet2D.py:1907-1930).
    double sum = 0.0;
    const volatile double *part = a.partial + (size_t)m * nstrips;
    for (int i = 0; i < nstrips; ++i) sum += part[i];
    const double resnorm = sqrt(sum);
    const int k = ctl->k;
    double init = ctl->init_norm;
    if (k == 0) { init = resnorm; ctl->init_norm = resnorm; }   // x_0 == wp: init_norm == resnorms[0]
    const double rel = resnorm / init;
    const bool conv = rel < a.rel_tol;                           // test precedes the solve (:1919)
    const int P = ctl->slot_p, Q = ctl->slot_q;
    if (conv || k + 1 >= a.max_its) {
        // converged: the un-updated iterate x_k is the answer; cap reached: x_{k+1} is
        const int fin = conv ? in_slot : out_slot;
        const int other = (fin == P) ? Q : P;   // fin == wp_slot (test passed at k == 0) leaves P, Q free
        a.snap_slot[(size_t)m * (a.num_steps + 1) + t + 1] = fin;
        a.its[(size_t)m * a.num_steps + t] = k + 1;              // == len(resnorms)
        a.relnorm[(size_t)m * a.num_steps + t] = rel;
        const int fresh = ctl->next_fresh;
        ctl->t = t + 1; ctl->k = 0;
        ctl->wp_slot = fin; ctl->in_slot = fin;
        ctl->slot_p = other; ctl->slot_q = fresh; ctl->next_fresh = fresh + 1;
        ctl->out_slot = other;
        if (t + 1 == a.num_steps) atomicAdd(&a.glob->done, 1);
    } else {
        ctl->k = k + 1;
        ctl->in_slot = out_slot;
        ctl->out_slot = (out_slot == P) ? Q : P;
    }
}

}  // namespace bfd

namespace bfd {

// natural (u|v, row-major) -> skewed strip layout, for every sample's slot 0; also resets control
__global__ void fom_import_kernel(const FomArgs a, const double *w0, size_t w0_stride) {
    const Ssl &L = a.ssl;
    const size_t per_state = L.state_doubles;
    const size_t total = per_state * a.batch;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
         i += (size_t)gridDim.x * blockDim.x) {
        const int m = (int)(i / per_state);
        size_t o = i - (size_t)m * per_state;
        const int strip = (int)(o / L.strip_stride);
        o -= (size_t)strip * L.strip_stride;
        const int s = (int)(o / (2 * L.W));
        const int rem = (int)(o - (size_t)s * 2 * L.W);
        const int comp = rem / L.W, cc = rem % L.W;
        const int r = s - cc / L.cpl, c = strip * L.W + cc;
        double val = 0.0;
        if (r >= 0 && r < L.ny && c < L.nx)
            val = w0[(size_t)m * w0_stride + (size_t)comp * L.nx * L.ny + (size_t)r * L.nx + c];
        a.pool[(size_t)m * a.pool_stride_mu + o + (size_t)strip * L.strip_stride] = val;
    }
}

__global__ void fom_reset_kernel(const FomArgs a) {
    const size_t n_edge = a.edge_stride_mu * a.batch;
    const double sentinel = __longlong_as_double((long long)kSentinelBits);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n_edge;
         i += (size_t)gridDim.x * blockDim.x)
        a.edge[i] = sentinel;
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g == 0) { a.glob->ticket = 0; a.glob->done = (a.num_steps == 0) ? a.batch : 0; a.glob->error = 0; }
    if (g < a.batch) {
        FomCtl c;
        c.t = 0; c.k = 0; c.in_slot = 0; c.wp_slot = 0; c.slot_p = 1; c.slot_q = 2; c.out_slot = 1;
        c.next_fresh = 3; c.arrive = 0; c.pad0 = 0; c.init_norm = 0.0; c.pad1[0] = c.pad1[1] = 0.0;
        a.ctl[g] = c;
        a.snap_slot[(size_t)g * (a.num_steps + 1)] = 0;
    }
}

// skewed strip layout -> caller's array.  One CTA (32 x 8 threads) moves a tile of 32 slot
// entries x 32 time levels through shared memory so that both sides are coalesced: reads run
// along a slot-row (lanes), writes along time (layout 1) or along x (layout 0).
__global__ void fom_export_kernel(const FomArgs a, int m, int t0, int t1, int layout, size_t ld, double *out) {
    const Ssl &L = a.ssl;
    BFD_SHARED_ARRAY(double, tile, [32][33]);
    const int strip = blockIdx.y, s = blockIdx.x;      // one slot-row (u and v) per CTA column
    const int nt = t1 - t0;
    const size_t N = (size_t)L.nx * L.ny;
    const int *slots = a.snap_slot + (size_t)m * (a.num_steps + 1);
    const double *pool_m = a.pool + (size_t)m * a.pool_stride_mu;
    for (int comp = 0; comp < 2; ++comp) {
        for (int cc0 = 0; cc0 < L.W; cc0 += 32) {
            for (int tb = 0; tb < nt; tb += 32) {
                // load: tile[time][entry]
                for (int ty = threadIdx.y; ty < 32; ty += blockDim.y) {
                    const int t = tb + ty;
                    if (t < nt) {
                        const size_t o = (size_t)slots[t0 + t] * L.state_doubles +
                                         ((size_t)(strip * L.slot_rows + s) * 2 + comp) * L.W + cc0 + threadIdx.x;
                        tile[ty][threadIdx.x] = pool_m[o];
                    }
                }
                __syncthreads();
                if (layout == 1) {
                    // write: for each entry (cell), 32 consecutive time levels
                    for (int e = threadIdx.y; e < 32; e += blockDim.y) {
                        const int cc = cc0 + e, r = s - cc / L.cpl, c = strip * L.W + cc;
                        const int t = tb + threadIdx.x;
                        if (r >= 0 && r < L.ny && c < L.nx && t < nt)
                            out[((size_t)comp * N + (size_t)r * L.nx + c) * ld + t] = tile[threadIdx.x][e];
                    }
                } else {
                    for (int ty = threadIdx.y; ty < 32; ty += blockDim.y) {
                        const int t = tb + ty;
                        const int cc = cc0 + threadIdx.x, r = s - cc / L.cpl, c = strip * L.W + cc;
                        if (r >= 0 && r < L.ny && c < L.nx && t < nt)
                            out[(size_t)t * 2 * N + (size_t)comp * N + (size_t)r * L.nx + c] = tile[ty][threadIdx.x];
                    }
                }
                __syncthreads();
            }
        }
    }
}

}  // namespace bfd

// ------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------
using namespace bfd;

extern "C" size_t bfd_fom_workspace_bytes(const bfd_fom_desc *d) {
    if (check_desc(d) != BFD_OK) return 0;
    return fom_layout(d).total;
}

extern "C" int bfd_fom_init(const bfd_fom_desc *d, void *ws, size_t ws_bytes, void *stream) {
    if (check_desc(d) != BFD_OK || !ws) return BFD_EINVAL;
    const FomLayout L = fom_layout(d);
    if (ws_bytes < L.total) return BFD_EWORKSPACE;
    const FomArgs a = fom_args(d, L, ws);
    BFD_LAUNCH(fom_reset_kernel, BFD_GRID(296), 256, 0, (cudaStream_t)stream, a);
    BFD_LAUNCHED();
    return BFD_OK;
}

extern "C" int bfd_fom_set_initial(const bfd_fom_desc *d, void *ws, const double *w0, size_t w0_stride,
                                   void *stream) {
    if (check_desc(d) != BFD_OK || !ws || !w0) return BFD_EINVAL;
    const FomLayout L = fom_layout(d);
    const FomArgs a = fom_args(d, L, ws);
    BFD_LAUNCH(fom_import_kernel, BFD_GRID(592), 256, 0, (cudaStream_t)stream, a, w0, w0_stride);
    BFD_LAUNCHED();
    return BFD_OK;
}

template <int CPL>
static int launch_sweep(const FomArgs &a, cudaStream_t st) {
    constexpr int WARPS = 4, PF = 4;
    const int items = a.batch * a.ssl.nstrips;
    const int grid = (items + WARPS - 1) / WARPS;
    auto kernel = fom_sweep_kernel<CPL, PF, WARPS>;
    BFD_LAUNCH_CORESIDENT(kernel, grid, WARPS * 32, 0, st, a);
    BFD_LAUNCHED();
    return BFD_OK;
}

extern "C" int bfd_fom_run(const bfd_fom_desc *d, void *ws, const bfd_fom_coef *coef, void *h_pinned,
                           void *stream) {
    if (check_desc(d) != BFD_OK || !ws || !coef || !h_pinned) return BFD_EINVAL;
    const FomLayout L = fom_layout(d);
    FomArgs a = fom_args(d, L, ws);
    a.cx = coef->cx; a.cxl = coef->cxl; a.cy = coef->cy; a.cyd = coef->cyd;
    a.src = coef->src; a.lbc = coef->lbc;
    cudaStream_t st = (cudaStream_t)stream;
    volatile int *h = (volatile int *)h_pinned;
    h[0] = 0; h[1] = 0; h[2] = 0;
    if (d->num_steps == 0) { h[0] = d->batch; return BFD_OK; }

    // Samples advance independently on the device; the host only keeps sweeps queued and polls
    // the "samples finished" counter one batch of launches behind, so the GPU never idles on it.
    // A sweep launched after everything finished exits at once.
    cudaEvent_t ev[2];
    BFD_CUDA(cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming));
    BFD_CUDA(cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming));
    int *h_poll = (int *)h_pinned + 4;   // two 2-int landing zones at h[4..5], h[6..7]
    const long long hard_cap = (long long)d->num_steps * d->max_its + 8;
    long long launched = 0;
    int rc = BFD_OK, pending = 0, slot = 0;
    const int chunk = 8;
    bool finished = false;
    while (!finished && rc == BFD_OK) {
        for (int i = 0; i < chunk && rc == BFD_OK; ++i) {
            rc = (d->cells_per_lane == 1) ? launch_sweep<1>(a, st) : launch_sweep<2>(a, st);
            ++launched;
        }
        if (rc != BFD_OK) break;
        cudaError_t e = cudaMemcpyAsync(h_poll + 2 * slot, &a.glob->done, 2 * sizeof(int),
                                        cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaEventRecord(ev[slot], st);
        if (e != cudaSuccess) { rc = record_cuda_error(e, "fom poll"); break; }
        if (pending) {   // look at the poll issued one chunk ago
            const int prev = slot ^ 1;
            e = cudaEventSynchronize(ev[prev]);
            if (e != cudaSuccess) { rc = record_cuda_error(e, "fom poll sync"); break; }
            if (h_poll[2 * prev] >= d->batch || h_poll[2 * prev + 1] != 0) finished = true;
        }
        pending = 1;
        slot ^= 1;
        if (launched > hard_cap + 2 * chunk) { rc = BFD_ENOTCONVERGED; break; }
    }
    cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess && rc == BFD_OK) rc = record_cuda_error(e, "fom run sync");
    cudaEventDestroy(ev[0]);
    cudaEventDestroy(ev[1]);
    if (rc == BFD_OK) {
        int fin[2] = {0, 0};
        e = cudaMemcpy(fin, &a.glob->done, sizeof(fin), cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) return record_cuda_error(e, "fom final status");
        h[0] = fin[0]; h[1] = fin[1]; h[2] = (int)launched;
        if (fin[1] != 0) rc = fin[1];
        else if (fin[0] < d->batch) rc = BFD_ENOTCONVERGED;
    }
    return rc;
}

extern "C" int bfd_fom_export(const bfd_fom_desc *d, const void *ws, int mu_index, int t0, int t1, int layout,
                              size_t ld, double *out, void *stream) {
    if (check_desc(d) != BFD_OK || !ws || !out) return BFD_EINVAL;
    if (mu_index < 0 || mu_index >= d->batch || t0 < 0 || t1 > d->num_steps + 1 || t0 >= t1) return BFD_EINVAL;
    if (layout != 0 && layout != 1) return BFD_EINVAL;
    if (layout == 1 && ld < (size_t)(t1 - t0)) return BFD_EINVAL;
    const FomLayout L = fom_layout(d);
    const FomArgs a = fom_args(d, L, const_cast<void *>(ws));
    dim3 grid(d->ny + kSkew, L.ssl.nstrips), block(32, 8);
    BFD_LAUNCH(fom_export_kernel, grid, block, 0, (cudaStream_t)stream, a, mu_index, t0, t1, layout, ld, out);
    BFD_LAUNCHED();
    return BFD_OK;
}

extern "C" int bfd_fom_diagnostics(const bfd_fom_desc *d, const void *ws, int32_t *its, double *relnorm,
                                   void *stream) {
    if (check_desc(d) != BFD_OK || !ws) return BFD_EINVAL;
    const FomLayout L = fom_layout(d);
    const FomArgs a = fom_args(d, L, const_cast<void *>(ws));
    const size_t n = (size_t)d->batch * d->num_steps;
    if (its) BFD_CUDA(cudaMemcpyAsync(its, a.its, n * sizeof(int), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    if (relnorm)
        BFD_CUDA(cudaMemcpyAsync(relnorm, a.relnorm, n * sizeof(double), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return BFD_OK;
}
