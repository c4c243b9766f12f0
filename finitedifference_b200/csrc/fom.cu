// Full-order model: batched trapezoidal time steps with all Newton sweeps of a step fused into
// one pass of a register wavefront.
//
// Replaces, for the FOM path of the reference (BurgersFD_CleanFine/hypernet2D.py):
//   inviscid_burgers_implicit2D :72-131, newton_raphson :1887-1933,
//   inviscid_burgers_res2D_alt :2588-2646, inviscid_burgers_exact_jac2D :2703-2732 and the
//   SuperLU spsolve at :1930.
//
// One Newton iteration J d = r, x <- x - d never builds J: with first-order upwind differences
// the Jacobian is 2x2-block lower triangular (a cell couples to itself, its left and its
// lower-row neighbour), so the solve is a block forward substitution
//     D_p d_p = r_p - A_L d_L - A_D d_D
// in which the cells of an anti-diagonal are independent.  A warp owns a strip of 32 columns
// and walks that wavefront down the rows: lane l is l rows behind lane 0, the left neighbour's
// values arrive by __shfl_up, the upper neighbour's are the lane's own previous step.
//
// Because EVERY dependency points up-wind, Newton iteration k+1 of a cell needs nothing but
// iteration k of the same three cells.  The kernel therefore runs the K solves of a time step
// as K "levels" of one software pipeline: at wavefront step s a lane works on level k of row
// s - lane - k, all levels in the same step being independent of one another (instruction-level
// parallelism for the FP64 pipe).  A time step costs one read of x_t (16 B/cell) and one write
// of x_{t+1} (16 B/cell) instead of 48 B per Newton sweep, and several time steps are chained
// inside one launch, so HBM and launch latency stop being the bound; the FP64 pipe is.
//
// Levels also accumulate ||r(x_k)||^2, k = 0..K, so newton_raphson's stopping rule is checked
// for exactly the iterates the reference would test.  K is speculated (the previous step's
// count); the last strip of a sample to finish a step validates the speculation on the device:
// "converged at K" accepts the step, "converged earlier" redoes it with the right K (strict
// mode), "not converged" continues it with further in-place sweeps.  Samples progress
// independently; the host only keeps launches queued.
//
// States live in the skewed strip layout (common.cuh) so the anti-diagonal a warp touches per
// step is one contiguous line of HBM.  Strips are chained left to right through
// self-validating records in global memory (wavefront.cuh).
#include "wavefront.cuh"
#include <stdlib.h>

namespace bfd {

constexpr int kKMax = BFD_FOM_KMAX;
constexpr int kNormSlots = kKMax + 1;

constexpr unsigned kGateStop = 1u << 16;   // BFD_FOM_MAX_CHUNK attempts fit below
enum StepStatus { kStepNone = 0, kStepOk = 1, kStepOkCap = 2, kStepEarly = 3, kStepNotConverged = 4 };

struct FomCtl {            // one per mu sample, 64 bytes
    int t;                 // time level to advance next (t -> t+1); == num_steps when finished
    int kbase;             // solves already applied to step t by earlier launches (continuation)
    int k_first;           // solves for the first step of the next launch
    int k_spec;            // solves speculated for the steps after it
    unsigned gate;         // kGateStop | highest attempt any strip of the sample has started
    unsigned arrive;       // strips of this sample that finished the current segment
    int redone;            // steps redone or continued so far (statistics)
    unsigned epoch;        // bumped by every commit: strips waiting to restart watch it
    double init_norm;      // ||r(x_0)|| of step t, kept for continuation passes
    unsigned exits;        // slab runs: strips (of all ranks) that left a launch, summed over launches
    unsigned under;        // consecutive steps that needed a continuation pass (speculation policy)
    double pad1[2];
};
static_assert(sizeof(FomCtl) == 64, "FomCtl must stay 64 bytes");

struct StepRec {           // verdict on one time step of one launch
    int status, kconv;
    double rel, init;
};

struct FomGlobal {         // one per workspace, 64 bytes
    unsigned ticket;       // virtual block id dispenser (wraps every launch)
    int done;              // samples that completed all steps
    int error;             // BFD_E* raised on the device
    int redone;            // total steps redone / continued
    int pad[12];
};

// Which part of the grid a workspace holds.  Whole grid: strip0 = 0, ns_global = local strips.
// Slab runs (one grid over several GPUs, bfd_fom_slab): this rank owns the 32-column strips
// [strip0, strip0 + ssl.nstrips); the per-sample control blocks, norms and verdicts of ALL strips
// live in rank 0's workspace ("home", peer-mapped), the records its first strip consumes in its own
// "halo" buffer, which the last strip of the rank on the left fills over NVLink.
struct SlabView {
    bool slab;
    int rank, world, strip0, ns_global, nx_local;
    void *home;
    double *next_halo;
};

struct FomLayout {         // byte offsets inside the workspace
    Ssl ssl;               // the strips held here (nx = local columns)
    int batch, num_steps, chunk;
    size_t off_global, off_ctl, off_its, off_relnorm, off_steprec, off_arrive, off_partial, off_trace, off_edge,
        off_halo, off_pool, total;
    size_t edge_stride_strip;  // doubles: one strip boundary, [chunk][ny][kRecDoubles]
    size_t edge_stride_mu;     // doubles
    size_t pool_stride_mu;     // doubles
};

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

static int desc_chunk(const bfd_fom_desc *d) { return d->chunk_steps > 0 ? d->chunk_steps : 50; }
static int desc_hint(const bfd_fom_desc *d) {
    int h = d->newton_hint > 0 ? d->newton_hint : 4;
    if (h > d->max_its) h = d->max_its;
    return h;
}

static int check_desc(const bfd_fom_desc *d) {
    if (!d || d->nx < 1 || d->ny < 1 || d->batch < 1 || d->num_steps < 0 || d->max_its < 1) return BFD_EINVAL;
    if (d->chunk_steps < 0 || d->chunk_steps > BFD_FOM_MAX_CHUNK) return BFD_EINVAL;
    if (d->newton_hint < 0 || d->newton_hint > kKMax) return BFD_EINVAL;
    if (!(d->rel_tol < 1.0)) return BFD_EINVAL;
    if (d->kernel != 0 && d->kernel != 1) return BFD_EINVAL;
    return BFD_OK;
}

static SlabView whole_grid(const bfd_fom_desc *d) {
    SlabView v;
    v.slab = false; v.rank = 0; v.world = 1; v.strip0 = 0;
    v.ns_global = make_ssl(d->nx, d->ny, 1).nstrips;
    v.nx_local = d->nx; v.home = nullptr; v.next_halo = nullptr;
    return v;
}

static int slab_view(const bfd_fom_desc *d, const bfd_fom_slab *sl, SlabView *v) {
    if (check_desc(d) != BFD_OK || !sl) return BFD_EINVAL;
    const int ns = make_ssl(d->nx, d->ny, 1).nstrips;
    if (sl->world < 1 || sl->rank < 0 || sl->rank >= sl->world) return BFD_EINVAL;
    if (sl->strip_begin < 0 || sl->strip_end > ns || sl->strip_begin >= sl->strip_end) return BFD_EINVAL;
    if ((sl->rank == 0) != (sl->strip_begin == 0) || (sl->rank == sl->world - 1) != (sl->strip_end == ns)) return BFD_EINVAL;
    v->slab = true; v->rank = sl->rank; v->world = sl->world; v->strip0 = sl->strip_begin; v->ns_global = ns;
    const int c1 = sl->strip_end * 32 < d->nx ? sl->strip_end * 32 : d->nx;
    v->nx_local = c1 - sl->strip_begin * 32;
    v->home = sl->home; v->next_halo = (double *)sl->next_halo;
    return BFD_OK;
}

// Control arrays are sized by the GLOBAL strip count and come first, so their offsets are the
// same in every rank's workspace (the home pointers of a slab run are home + these offsets).
static FomLayout fom_layout(const bfd_fom_desc *d, const SlabView &v) {
    FomLayout L;
    L.ssl = make_ssl(v.nx_local, d->ny, 1);
    L.batch = d->batch;
    L.num_steps = d->num_steps;
    L.chunk = desc_chunk(d);
    const size_t B = (size_t)d->batch, S = (size_t)L.chunk, ns = (size_t)v.ns_global, nl = (size_t)L.ssl.nstrips;
    size_t o = 0;
    L.off_global = o;   o = align_up(o + sizeof(FomGlobal), 256);
    L.off_ctl = o;      o = align_up(o + sizeof(FomCtl) * B, 256);
    L.off_its = o;      o = align_up(o + sizeof(int) * B * d->num_steps, 256);
    L.off_relnorm = o;  o = align_up(o + sizeof(double) * B * d->num_steps, 256);
    L.off_steprec = o;  o = align_up(o + sizeof(StepRec) * B * S, 256);
    L.off_arrive = o;   o = align_up(o + sizeof(unsigned) * B * S, 256);
    L.off_partial = o;  o = align_up(o + sizeof(double) * B * S * ns * kNormSlots, 256);
    L.off_trace = o;    o = align_up(o + sizeof(unsigned long long) * B * S * ns * 3, 256);
    L.edge_stride_strip = S * (size_t)d->ny * kRecDoubles;
    L.edge_stride_mu = L.edge_stride_strip * (nl > 1 ? nl - 1 : 0);
    L.off_edge = o;     o = align_up(o + sizeof(double) * L.edge_stride_mu * B, 256);
    L.off_halo = o;     o = align_up(o + sizeof(double) * L.edge_stride_strip * B * (v.strip0 > 0 ? 1 : 0), 256);
    L.pool_stride_mu = L.ssl.state_doubles * (size_t)(d->num_steps + 1);
    L.off_pool = o;     o = align_up(o + sizeof(double) * L.pool_stride_mu * B, 256);
    L.total = o;
    return L;
}

// Arguments of the kernels (passed by value).
struct FomArgs {
    Ssl ssl;
    int batch, num_steps, max_its, chunk, strict, uniform_y, hint;
    double rel_tol;
    const double *cx, *cxl, *cy, *cyd, *src, *lbc;
    FomGlobal *glob;
    FomCtl *ctl;
    int *its;           // [batch][num_steps]
    double *relnorm;    // [batch][num_steps]
    StepRec *steprec;   // [batch][chunk]
    unsigned *arrive;   // [batch][chunk]
    double *partial;    // [batch][chunk][nstrips][kNormSlots]
    unsigned long long *trace;   // [batch][chunk][nstrips][3] globaltimer at the start / end of a strip's time step, hand-off misses << 32 | spins
    double *edge;       // [batch][nstrips-1][chunk][ny][kRecDoubles]
    double *pool;       // [batch][num_steps+1][state]   snapshot t of sample m in slot t
    size_t edge_stride_strip, edge_stride_mu, pool_stride_mu;
    // slab runs (whole grid: strip0 = 0, ns_global = ssl.nstrips, the rest unused)
    int strip0, ns_global, launch_index, slab;
    int variant;        // 0: fom_chunk_kernel (levels split over warps), 1: fom_tile_kernel (warp per strip and time step)
    int kcap;           // Newton solves the variant fuses into one pass at most
    unsigned *ticket;   // virtual block id dispenser of THIS rank (glob may be rank 0's)
    double *halo;       // [batch][chunk][ny][kRecDoubles] records consumed by this rank's first strip
    double *next_halo;  // the same buffer of the rank on the right (peer-mapped), or null
};

static FomArgs fom_args(const bfd_fom_desc *d, const FomLayout &L, void *ws, const SlabView &v) {
    char *b = (char *)ws;                                     // pool, hand-off records, ticket, trace: local
    char *h = (v.slab && v.home) ? (char *)v.home : b;        // control blocks, norms, verdicts: rank 0's
    FomArgs a;
    a.ssl = L.ssl;
    a.batch = d->batch; a.num_steps = d->num_steps; a.max_its = d->max_its; a.chunk = L.chunk;
    a.strict = d->strict; a.uniform_y = d->uniform_y; a.hint = desc_hint(d);
    a.rel_tol = d->rel_tol;
    a.cx = a.cxl = a.cy = a.cyd = a.src = a.lbc = nullptr;
    a.glob = (FomGlobal *)(h + L.off_global);
    a.ctl = (FomCtl *)(h + L.off_ctl);
    a.its = (int *)(h + L.off_its);
    a.relnorm = (double *)(h + L.off_relnorm);
    a.steprec = (StepRec *)(h + L.off_steprec);
    a.arrive = (unsigned *)(h + L.off_arrive);
    a.partial = (double *)(h + L.off_partial);
    a.ticket = &((FomGlobal *)(b + L.off_global))->ticket;
    a.strip0 = v.strip0; a.ns_global = v.ns_global; a.launch_index = 0; a.slab = v.slab ? 1 : 0;
    a.variant = d->kernel; a.kcap = d->kernel == 1 ? 4 : kKMax;     // kTileKMax (fom_tile.cuh)
    a.halo = v.strip0 > 0 ? (double *)(b + L.off_halo) : nullptr;
    a.next_halo = v.next_halo;
    a.trace = (unsigned long long *)(b + L.off_trace);
    a.edge = (double *)(b + L.off_edge);
    a.pool = (double *)(b + L.off_pool);
    a.edge_stride_strip = L.edge_stride_strip;
    a.edge_stride_mu = L.edge_stride_mu;
    a.pool_stride_mu = L.pool_stride_mu;
    return a;
}

}  // namespace bfd

namespace bfd {

#ifndef BFD_MINB
#define BFD_MINB 4
#endif
#ifndef BFD_WARPS
#define BFD_WARPS 4
#endif
constexpr int kWarpsPerStrip = BFD_WARPS;     // warps sharing the Newton levels of one strip

// Shared memory of one CTA (byte sizes).  Everything a lane exchanges is a 16-byte pair, laid
// out [pair][lane] so that 128-bit accesses of a quarter warp are conflict-free.
constexpr int kPairRow = 32 * 16;              // one pair for each lane
constexpr int kHandBytes = 2 * 3 * kPairRow;   // level hand-off between two warps: [2 buffers][(xu,xv),(Au,Av),(cy,cyd)][lane]
constexpr int kHandIfaces = kWarpsPerStrip > 1 ? kWarpsPerStrip - 1 : 1;
constexpr int kXchgLevels = kKMax + 2;         // levels 0..K, plus (Fp, Gp) of continuation passes
constexpr int kXRow = 33 * 16;                  // exchange row: slot 0 = halo (last column of the strip on the left), slot l+1 = lane l
constexpr int kXchgBytes = kXchgLevels * 2 * 2 * kXRow;      // left-neighbour exchange: [level][2 buffers][(Fh,Gh),(pu,pm)][slot]
constexpr int kPfDepth = 6;                    // rows of level-0 input in flight (cp.async ring)
constexpr int kPfPending = kPfDepth - 3;       // commit groups that may be in flight after a step's wait: rows <= s + 2 have landed
                                               // after step s, i.e. the row used in step s + 1 and the one it loads ahead
constexpr int kPfValues = 7;                   // u, v, up, vp (CONT), lbc (inlet lane), cy, cyd (non-uniform grids)
constexpr int kPfBytes = kPfDepth * kPfValues * 32 * 8;

template <bool B> struct BoolTag { static constexpr bool value = B; };

// What one warp needs to run its levels of one time step of its strip.
struct StepIo {
    const double *xin;    // level-0 iterate: snapshot t, or the partly converged x of slot t+1 (CONT)
    const double *xwp;    // snapshot t
    double *xout;         // slot t+1 (all three already offset to this strip and lane)
    double *edge_in;      // hand-off records of the strip on the left for this time step
    double *edge_out;     // ... and the ones this strip produces
    saddr ring_in;        // staging ring of the left strip's records (shared-window address)
    saddr hand;           // level hand-off written by this warp (the previous warp's is kHandBytes below)
    saddr pf;             // prefetch ring of the level-0 input [kPfDepth][kPfValues][32]
    saddr xchg;           // left-neighbour exchange
    double *partial;      // [kNormSlots] squared residual norms of this (sample, step, strip)
    const double *lbc;    // [ny] inlet term of this sample
    const double *cy, *cyd;
    int ny, strip, nstrips, lane;
    double cx, cxl, src;  // column constants of this lane
    bool colok, inlet;
    bool in_remote, out_remote;   // slab runs: the hand-off on that side crosses to another GPU
    int *error;
    unsigned long long *stats;   // hand-off wait statistics of this (sample, step, strip)
};

struct Level {            // what a cell keeps for its lower neighbour (D) at one level
    double Gh, Hh;        // current-iterate flux parts uv/2, v^2/2
    double pm, pvv;       // (v*du + u*dv) / 2, v*dv
};

// Two pieces of the CTA's data movement are duties that any warp can carry: they touch nothing but
// shared memory that the per-step barrier already orders and their own running addresses.  With four
// compute warps they ride on the two warps that hold one solve level each and would otherwise wait
// ~250 cycles per step at the barrier (r02: the data-movement warp was the last to arrive, 7 % barrier
// stall against 42-45 % for those two), in the stall slots of their FP64 chains.
enum Duty { kDutyNone = 0, kDutyFeed = 1, kDutyGive = 2 };
#ifndef BFD_SPLIT
#define BFD_SPLIT 0
#endif
#ifndef BFD_UNROLL
#define BFD_UNROLL 2
#endif
constexpr int kMainUnroll = BFD_UNROLL;       // wavefront steps per trip of the main loop
constexpr bool kSplitDuties = BFD_SPLIT && (kWarpsPerStrip == 4);

// kDutyFeed: keeps the level-0 input ring kPfDepth-1 rows ahead with cp.async, one row and one commit
// group per step: "all but the newest kPfPending groups have landed" = "the rows of the next two steps are there".
template <bool CONT, bool UNI>
struct InputFeed {
    saddr pf_lane, pf_end, pf_d;
    const double *src, *srcp, *lbc_p, *cy, *cyd;
    int pf_row, ny, last_row, lane;
    bool strip0;
    __device__ __forceinline__ explicit InputFeed(const StepIo &io) {
        lane = io.lane; ny = keep_reg(io.ny); last_row = keep_reg(ny + kSkew - 1);
        pf_lane = keep_reg(io.pf + (saddr)(lane * 8));
        pf_end = keep_reg(pf_lane + (saddr)(kPfDepth * kPfValues * 256));
        pf_d = pf_lane;
        src = io.xin; srcp = io.xwp; lbc_p = io.lbc; cy = io.cy; cyd = io.cyd;
        pf_row = 0;
        strip0 = io.strip == 0;
    }
    __device__ __forceinline__ void issue() {
        constexpr int W = 32;
        cp_async8(pf_d, src);
        cp_async8(pf_d + 256, src + W);
        if (CONT) { cp_async8(pf_d + 512, srcp); cp_async8(pf_d + 768, srcp + W); }
        if (strip0) {                                        // warp-uniform: only the inlet strip carries lbc
            if (lane == 0) cp_async8(pf_d + 1024, lbc_p);
            if (pf_row < ny - 1) ++lbc_p;
        }
        if (!UNI) {
            const int r = pf_row < last_row ? pf_row : last_row;
            const int rcy = r - lane < 0 ? 0 : (r - lane >= ny ? ny - 1 : r - lane);   // grid row of this lane's cell
            cp_async8(pf_d + 1280, cy + rcy);
            cp_async8(pf_d + 1536, cyd + rcy);
        }
        if (pf_row < last_row) { src += 2 * W; if (CONT) srcp += 2 * W; }
        ++pf_row;
        pf_d += (saddr)(kPfValues * 256);
        if (pf_d == pf_end) pf_d = pf_lane;
    }
    // the same for a row that is neither the last row of the state nor (inlet strip) past the last grid row
    __device__ __forceinline__ void step_main() {
        constexpr int W = 32;
        static_assert(UNI || !UNI, "");
        cp_async8(pf_d, src);
        cp_async8(pf_d + 256, src + W);
        if (CONT) { cp_async8(pf_d + 512, srcp); cp_async8(pf_d + 768, srcp + W); }
        if (strip0) { if (lane == 0) cp_async8(pf_d + 1024, lbc_p); ++lbc_p; }
        src += 2 * W; if (CONT) srcp += 2 * W;
        ++pf_row;
        pf_d += (saddr)(kPfValues * 256);
        if (pf_d == pf_end) pf_d = pf_lane;
        cp_async_commit();
        cp_async_wait_pending<kPfPending>();
    }
    __device__ __forceinline__ void prologue() {             // before the barrier that opens step 0
#pragma unroll
        for (int i = 0; i < kPfDepth - 1; ++i) { issue(); cp_async_commit(); }
        cp_async_wait_pending<kPfPending>();
    }
    __device__ __forceinline__ void step() {                 // row s + kPfDepth - 1: refills the slot read one step ago
        issue();
        cp_async_commit();
        cp_async_wait_pending<kPfPending>();
    }
};

// kDutyGive: sends what lane 31 published in the PREVIOUS step (slot 32 of the exchange rows) to the strip
// on the right: lane p moves piece p of the row records, one 16-byte piece per lane and step, straight to
// the records in global (or peer) memory; the row a lane sends lags by the Newton level of its piece.
template <int K, bool CONT>
struct EdgeGive {
    static constexpr int NPR = 2 * K + 1 + (CONT ? 1 : 0);
    saddr x_prev, x_sum;
    double *gout;
    int ro, ny;
    bool give, out_remote;
    __device__ __forceinline__ explicit EdgeGive(const StepIo &io) {
        const int lane = io.lane;
        const bool special = CONT && lane == 2 * K + 1;
        const int lvl = special ? 0 : (lane >> 1);
        ny = io.ny;
        give = lane < NPR && io.strip < io.nstrips - 1;
        out_remote = io.out_remote;
        const saddr p_x = io.xchg + (saddr)((special ? K + 1 : (lane >> 1)) * 4 * kXRow + (special ? 0 : (lane & 1)) * kXRow);
        x_sum = 2 * p_x + (saddr)(2 * kXRow);                // the two exchange buffers alternate: next = x_sum - current
        x_prev = p_x;                                        // buffer of step -2 (even)
        ro = -2 - kSkew - lvl;                               // row sent by the call for step s: s - 1 - 31 - level
        gout = io.edge_out + 2 * lane + (ptrdiff_t)ro * kRecDoubles;
    }
    __device__ __forceinline__ void step() {                 // once before step 0, once per step, once after the last
        if (give && (unsigned)ro < (unsigned)ny) {
            double a, b;
            lds2(x_prev + 32 * 16, a, b);
            if (out_remote) st_sys2(gout, a, b); else st_volatile2(gout, a, b);
        }
        ++ro; gout += kRecDoubles;
        x_prev = x_sum - x_prev;
    }
    __device__ __forceinline__ void step_main() {            // the row sent is a grid row for every lane
        if (give) {
            double a, b;
            lds2(x_prev + 32 * 16, a, b);
            if (out_remote) st_sys2(gout, a, b); else st_volatile2(gout, a, b);
        }
        ++ro; gout += kRecDoubles;
        x_prev = x_sum - x_prev;
    }
};

// Levels LO..HI-1 of one trapezoidal time step of one strip.  A time step is K fused Newton
// solves (levels 0..K-1) plus the residual of x_K (level K); the levels are split between the
// warps of the CTA, which pass the iterate from level HI-1 to level HI through shared memory
// and meet at one CTA barrier per wavefront step.
//   CONT  continuation pass: the level-0 iterate differs from the previous snapshot
//   UNI   cy / cyd are the same for every row (kept in registers)
// Residual in the form used here (hypernet2D.py:2619-2643 regrouped): with Fh = u^2/2,
// Gh = uv/2, Hh = v^2/2 of the iterate and Fp, Gp, Hp of the previous snapshot,
//   ru = [u + cx Fh + cy Gh - cxl Fh_L - cyd Gh_D] + Au,  Au = -up + cx Fp + cy Gp - cxl Fp_L - cyd Gp_D - src - lbc
//   rv = [v + cy Hh + cx Gh - cyd Hh_D - cxl Gh_L] + Av,  Av = -vp + cy Hp + cx Gp - cyd Hp_D - cxl Gp_L
// Au, Av do not change between Newton levels and ride along the level pipeline.
//
// Neighbours: the upper cell (r-1, c) is the lane's own previous step (registers).  The left
// cell (r, c-1) is lane-1's previous step: every lane publishes (Fh, Gh) and (u du, v du + u dv)
// of each level in a double-buffered shared-memory exchange and reads its left neighbour's one
// step later; lane 0 reads the same pairs from the hand-off record of the strip on its left
// instead (only the ADDRESS differs), and lane 31 also files them in the record for the strip
// on its right.
template <int K, int LO, int HI, bool CONT, bool UNI, int DUTY>
__device__ __forceinline__ void sweep_part(const StepIo &io) {
    constexpr bool FIRST = (LO == 0), LAST = (HI == K + 1);
    constexpr int W = 32;
    const int lane = io.lane, ny = io.ny;
    const int last_row = ny + kSkew - 1;   // last slot-row holding cells
    const int nsteps = ny + kSkew + K;

    const bool colok = io.colok;

    // The 0.5 factors of the Jacobian ride on the iterate (hu = u/2, hv = v/2) and on the published pair
    // (hpm = (v du + u dv)/2), exact power-of-two rescalings: the same bits as with halved coefficients, and
    // four coefficient registers (cx, cxl, cy, cyd) instead of eight -- ptxas recomputed the halved ones with
    // four DMULs per level and step rather than hold them.
    const double cx = io.cx, cxl = io.cxl, ncxl = -io.cxl;
    const double src = io.src;
    double cyU = 0.0, cydU = 0.0;
    if (UNI) { cyU = __ldg(io.cy + (ny > 1 ? 1 : 0)); cydU = __ldg(io.cyd + (ny > 1 ? 1 : 0)); }

    // level-0 input: rows land in the shared-memory ring one step ahead (dma_part)
    const saddr pf_lane = keep_reg(io.pf + (saddr)(lane * 8));
    const saddr pf_end = keep_reg(pf_lane + (saddr)(kPfDepth * kPfValues * 256));
    saddr pf_rd = pf_lane;                  // slot of the row read this step (running, no modulo)

    Level prev[K + 1];        // own cell of the previous step at each level = upper neighbour now
    double qGp = 0.0, qHp = 0.0;   // previous-snapshot flux parts of the previous row (CONT)
#pragma unroll
    for (int k = 0; k <= K; ++k) prev[k] = Level{0.0, 0.0, 0.0, 0.0};
    double xu[K + 2], xv[K + 2], Au[K + 2], Av[K + 2], Cy[K + 2], Cyd[K + 2], nrm[K + 1];
#pragma unroll
    for (int k = 0; k <= K + 1; ++k) { xu[k] = xv[k] = Au[k] = Av[k] = 0.0; Cy[k] = cyU; Cyd[k] = cydU; }
#pragma unroll
    for (int k = 0; k <= K; ++k) nrm[k] = 0.0;
    const saddr hand_in = io.hand - (saddr)kHandBytes + (saddr)(lane * 16);   // written by the previous warp
    const saddr hand_out = io.hand + (saddr)(lane * 16);
    const saddr xc_own = io.xchg + (saddr)((lane + 1) * 16);   // this lane's pairs (slot lane + 1)
    const saddr xc_left = io.xchg + (saddr)(lane * 16);        // the left column's: lane - 1, or the halo slot for lane 0
    auto xc_level = [](int k, int buf) { return (saddr)((k * 2 + buf) * 2 * kXRow); };
    InputFeed<CONT, UNI> feed(io);
    EdgeGive<K, CONT> edge(io);
    // level-0 input of the NEXT step, read from the ring at the end of a step: its shared-memory latency is
    // spent beside the tail of this step's chain and the barrier instead of at the head of the next chain
    struct { double u, v, up, vp, lb, cy, cyd; } nx = {0.0, 0.0, 0.0, 0.0, 0.0, cyU, cydU};
    auto load_row = [&]() {
        const saddr d = pf_rd;
        pf_rd += (saddr)(kPfValues * 256);
        if (pf_rd == pf_end) pf_rd = pf_lane;
        nx.u = lds(d); nx.v = lds(d + 256);
        if (CONT) { nx.up = lds(d + 512); nx.vp = lds(d + 768); }
        if (io.inlet) nx.lb = lds(d + 1024);
        if (!UNI) { nx.cy = lds(d + 1280); nx.cyd = lds(d + 1536); }
    };

    // One wavefront step.  EP ("edge phase"): some lane's row is outside the grid at some level
    // (first 32+K and last 31+K steps); the main phase runs without row masks.
    auto body = [&](const int s, auto ep) {
        constexpr bool EP = decltype(ep)::value;
        const int rbuf = (s + 1) & 1, wbuf = s & 1;      // exchange buffers: written last step / this step
        double up = 0.0, vp = 0.0, lb = 0.0;
        if (FIRST) {
            // the row of this step was loaded before the barrier that opened it (load_row)
            xu[0] = nx.u; xv[0] = nx.v;
            up = CONT ? nx.up : xu[0]; vp = CONT ? nx.vp : xv[0];
            lb = nx.lb;
            if (!UNI) { Cy[0] = nx.cy; Cyd[0] = nx.cyd; }
        } else {
            // x_LO and the level-independent terms of row s - lane - LO, from the previous warp
            const saddr h = hand_in + (saddr)(rbuf * 3 * kPairRow);
            lds2(h, xu[LO], xv[LO]);
            lds2(h + kPairRow, Au[LO], Av[LO]);
            if (!UNI) lds2(h + 2 * kPairRow, Cy[LO], Cyd[LO]);
        }
        if (DUTY == kDutyFeed) feed.step();
        if (DUTY == kDutyGive) edge.step();

        // levels in descending order: level k+1 consumes xu[k+1], Au[k+1] before level k
        // replaces them with the values for the next step
#pragma unroll
        for (int k = HI - 1; k >= LO; --k) {
            const int rk = s - lane - k;
            const bool rowok = EP ? (unsigned)rk < (unsigned)ny : true;
            const bool valid = rowok && colok;
            const bool top = EP ? (rk <= 0) : false;   // no upper neighbour (row -1 is not a cell)
            const double cy = UNI ? cyU : Cy[k], cyd = UNI ? cydU : Cyd[k];
            const double ncyd = -cyd;

            // left neighbour (r, c-1): the pairs slot `lane` received in the previous step -- from
            // lane - 1, or (lane 0) from the strip on the left through the data-movement warp
            const saddr la = xc_left + xc_level(k, rbuf);
            double LF, LG, Lpu = 0.0, Lpm = 0.0;
            lds2(la, LF, LG);
            if (k < K) lds2(la + kXRow, Lpu, Lpm);

            if (k == 0) {
                // previous-snapshot parts and the level-independent terms Au, Av
                const double hup = 0.5 * up, hvp = 0.5 * vp;
                const double Fp = hup * up, Gp = hup * vp, Hp = hvp * vp;
                double DGp = CONT ? qGp : prev[0].Gh, DHp = CONT ? qHp : prev[0].Hh;
                if (EP) { DGp = top ? 0.0 : DGp; DHp = top ? 0.0 : DHp; }
                double lFp = LF, lGp = LG;
                if (CONT) {
                    lds2(xc_left + xc_level(K + 1, rbuf), lFp, lGp);
                }
                double au = fma(cx, Fp, -up);
                au = fma(cy, Gp, au); au = fma(ncxl, lFp, au); au = fma(ncyd, DGp, au);
                au = (au - src) - lb;
                double av = fma(cy, Hp, -vp);
                av = fma(cx, Gp, av); av = fma(ncyd, DHp, av); av = fma(ncxl, lGp, av);
                Au[0] = au; Av[0] = av;
                if (CONT) {
                    qGp = Gp; qHp = Hp;
                    sts2(xc_own + xc_level(K + 1, wbuf), Fp, Gp);
                }
            }

            const double u = xu[k], v = xv[k];
            const double hu = 0.5 * u, hv = 0.5 * v;
            const double Fh = hu * u, Gh = hu * v, Hh = hv * v;
            double DG = prev[k].Gh, DH = prev[k].Hh;
            if (EP) { DG = top ? 0.0 : DG; DH = top ? 0.0 : DH; }
            double ru = Au[k] + u;
            ru = fma(cx, Fh, ru); ru = fma(cy, Gh, ru); ru = fma(ncxl, LF, ru); ru = fma(ncyd, DG, ru);
            double rv = Av[k] + v;
            rv = fma(cy, Hh, rv); rv = fma(cx, Gh, rv); rv = fma(ncyd, DH, rv); rv = fma(ncxl, LG, rv);
            const double r2 = fma(rv, rv, ru * ru);
            // main phase: every row is a cell; padding columns (last strip) are dropped after the loop
            nrm[k] += EP ? (valid ? r2 : 0.0) : r2;
            sts2(xc_own + xc_level(k, wbuf), Fh, Gh);
            prev[k].Gh = Gh; prev[k].Hh = Hh;

            if (k < K) {
                double Dpm = prev[k].pm, Dpvv = prev[k].pvv;
                if (EP) { Dpm = top ? 0.0 : Dpm; Dpvv = top ? 0.0 : Dpvv; }
                const double bu_ = fma(cyd, Dpm, fma(cxl, Lpu, ru));       // Dpm, Lpm hold the halved pair
                const double bv_ = fma(cyd, Dpvv, fma(cxl, Lpm, rv));
                const double d00 = fma(cx, u, fma(cy, hv, 1.0));
                const double d11 = fma(cy, v, fma(cx, hu, 1.0));
                const double d01 = cy * hu, d10 = cx * hv;
                const double inv = rcp_fast(fma(d00, d11, -(d01 * d10)));
                const double du = fma(d11, bu_, -(d01 * bv_)) * inv;
                const double dv = fma(d00, bv_, -(d10 * bu_)) * inv;
                const double pu = u * du, pm = fma(hv, du, hu * dv);       // (v du + u dv) / 2
                prev[k].pm = pm; prev[k].pvv = v * dv;
                sts2(xc_own + xc_level(k, wbuf) + kXRow, pu, pm);
                xu[k + 1] = u - du; xv[k + 1] = v - dv;
                Au[k + 1] = Au[k]; Av[k + 1] = Av[k];
                if (!UNI) { Cy[k + 1] = Cy[k]; Cyd[k + 1] = Cyd[k]; }
                if (k == K - 1) {
                    // x_K of row rk: slot-row (rk + lane) of the output state
                    const int so = s - (K - 1);
#ifdef BFD_X_NOSTORE
                    if (so == 0) {
#else
                    if (!EP || (so >= 0 && so <= last_row)) {
#endif
                        double *o = io.xout + (size_t)so * (2 * W);
                        // padding columns of the last strip evolve as a phantom continuation of the
                        // grid: finite, and never read by a real cell (information only moves right)
                        o[0] = EP ? (valid ? xu[K] : 0.0) : xu[K];
                        o[W] = EP ? (valid ? xv[K] : 0.0) : xv[K];
                    }
                }
            }
        }
        if (!LAST) {
            const saddr h = hand_out + (saddr)(wbuf * 3 * kPairRow);
            sts2(h, xu[HI], xv[HI]);
            sts2(h + kPairRow, Au[HI], Av[HI]);
            if (!UNI) sts2(h + 2 * kPairRow, Cy[HI], Cyd[HI]);
        }
        if (FIRST) load_row();               // row s + 1: landed before the barrier that opened this step
        __syncthreads();
    };

    if (DUTY == kDutyFeed) feed.prologue();
    if (DUTY == kDutyGive) edge.step();
    __syncthreads();      // rows 0, 1 and the first hand-off block are in shared memory (dma_part / the feed duty)
    if (FIRST) load_row();
    // head (rows entering the strip), main (0 < row < ny for every lane and level), tail; the
    // two edge phases share one copy of the masked body (code size: instruction cache)
    const int head_end = nsteps < kSkew + K + 1 ? nsteps : kSkew + K + 1;
    const int main_end = ny > head_end ? ny : head_end;
    int s = 0;
#pragma unroll 1
    for (int phase = 0; phase < 2; ++phase) {
        const int ep_end = phase == 0 ? head_end : nsteps;
#pragma unroll 1
        for (; s < ep_end; ++s) body(s, BoolTag<true>{});
        if (phase == 0) {
#pragma unroll kMainUnroll
            for (; s < main_end; ++s) body(s, BoolTag<false>{});
        }
    }

    if (DUTY == kDutyGive) edge.step();          // what lane 31 published in the last step
    // ||r(x_k)||^2 of this strip, fixed reduction order (deterministic iteration counts)
#pragma unroll
    for (int k = LO; k < HI; ++k) {
        double x = colok ? nrm[k] : 0.0;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) x += __shfl_xor_sync(0xffffffffu, x, off);
        if (lane == 0) io.partial[k] = x;
    }
}

// A warp with no level in this pass keeps the CTA's barrier count and carries its duty.
template <int K, bool CONT, bool UNI, int DUTY>
__device__ __forceinline__ void idle_part(const StepIo &io) {
    const int nsteps = io.ny + kSkew + K;
    InputFeed<CONT, UNI> feed(io);
    EdgeGive<K, CONT> edge(io);
    if (DUTY == kDutyFeed) feed.prologue();
    if (DUTY == kDutyGive) edge.step();
    __syncthreads();
#pragma unroll 1
    for (int s = 0; s < nsteps; ++s) {
        if (DUTY == kDutyFeed) feed.step();
        if (DUTY == kDutyGive) edge.step();
        __syncthreads();
    }
    if (DUTY == kDutyGive) edge.step();
}

// A piece that had not arrived when its row was copied into the ring: polled from global (or peer) memory.
// Out of line: the per-step loop of the data-movement warp stays one run of instructions in the
// instruction cache.  Inside the per-step loop a plain counter bounds the wait (the producer is running).
struct PolledPiece { double a, b; unsigned spins; int ok; };
__device__ __noinline__ PolledPiece poll_piece(const double *gin, bool in_remote) {
    PolledPiece p{0.0, 0.0, 0u, 1};
    for (;;) {
        if (in_remote) ld_sys2(gin, p.a, p.b); else ld_volatile2(gin, p.a, p.b);
        if (!(is_sentinel(p.a) || is_sentinel(p.b))) return p;
        if (++p.spins > kSpinLimit) { p.a = 0.0; p.b = 0.0; p.ok = 0; return p; }
        if (p.spins > 2) spin_pause();
    }
}

// The data-movement warp of the CTA: keeps the level-0 input ring kPfDepth-1 rows ahead with
// cp.async, brings the hand-off blocks of the strip on the left into the shared-memory ring
// ahead of time (spinning there, off the compute warps' critical path, if the producer is late),
// and every step
//   - drops the pairs lane 0 needs in the NEXT step into the halo slot of the exchange rows, so
//     that the compute warps treat all 32 lanes alike (no address select, no branch), and
//   - sends what lane 31 published in the PREVIOUS step to the strip on the right: one 16-byte
//     piece per lane straight to the records in global (or peer) memory.
// It meets the compute warps at the same per-step barrier.
template <int K, bool CONT, bool UNI>
__device__ __forceinline__ void dma_part(const StepIo &io) {
    constexpr int NPR = 2 * K + 1 + (CONT ? 1 : 0);
    static_assert(NPR <= 32, "one lane per piece of a record row");
    const int lane = io.lane, ny = keep_reg(io.ny);
    const int nsteps = ny + kSkew + K;
    int err = 0;
    unsigned misses = 0, spins_total = 0;
    const double armed = keep_reg(sentinel_value());         // what a consumed piece is overwritten with

    // ---- strip-to-strip hand-off: lane p moves piece p of the row records -------------------
    // Every lane runs its own stream of rows, one per step, offset by the Newton level its piece
    // belongs to (the previous-time-level pair of continuation passes travels with level 0):
    //   in : the halo slot of the exchange rows <- what lane 0 of the compute warps reads in the
    //        next step.  The piece comes from the shared-memory ring, where it was copied kAhead
    //        rows earlier (cp.async, in the per-step commit groups), and is read from the ring one
    //        step before it is needed so that no load is waited for inside a step; a piece that had
    //        not arrived when it was copied is polled from global memory.  Consumed pieces are
    //        re-armed (sentinel) for the next launch.
    //   out: the records of this strip <- what lane 31 (slot 32) published in the previous step.
    const bool special = CONT && lane == 2 * K + 1;
    const int lvl = special ? 0 : (lane >> 1);
    const bool take = lane < NPR && io.strip > 0;
    const bool in_remote = io.in_remote;
    const saddr p_x = io.xchg + (saddr)((special ? K + 1 : (lane >> 1)) * 4 * kXRow + (special ? 0 : (lane & 1)) * kXRow);
    const saddr x_sum = 2 * p_x + (saddr)(2 * kXRow);        // the two exchange buffers alternate: next = x_sum - current
    saddr x_cur = p_x + (saddr)(2 * kXRow);                  // buffer of step -1 (odd)
    InputFeed<CONT, UNI> feed(io);                           // the two duties, when no compute warp carries them
    EdgeGive<K, CONT> edge(io);
    const saddr ring0 = keep_reg(io.ring_in + (saddr)(lane * 16)), ring_end = keep_reg(ring0 + (saddr)(kInRing * kRecBytes));
    int rc = -lvl;                                           // row consumed by the call for step s: s + 1 - level
    double *gin = io.edge_in + 2 * lane + (ptrdiff_t)rc * kRecDoubles;
    saddr sl = ring0 + (saddr)(((rc + 1) & (kInRing - 1)) * kRecBytes);          // slot of the row loaded early
    saddr pl = ring0 + (saddr)(((rc + kAhead) & (kInRing - 1)) * kRecBytes);     // slot of the row copied now
    double va = 0.0, vb = 0.0;                               // the piece consumed by the next call

    if (io.strip > 0) {
        // Keep a distance to the producer: do not start the time step before it is kStartRow rows
        // into it.  Strips run at the same pace, so the distance stays and the copies find their
        // data; a strip that starts as soon as row 0 exists would stay that close for ever and
        // wait for a global-memory round trip in every step.
        const int rw = kStartRow < ny ? kStartRow : ny - 1;
        const double *gw = io.edge_in + (size_t)rw * kRecDoubles + 2 * ((NPR - 1) & ~1);   // last Newton level: written last
        SpinClock spins;
        for (;;) {
            double a = 0.0, b = 0.0;
            if (lane == 0) { if (in_remote) ld_sys2(gw, a, b); else ld_volatile2(gw, a, b); }
            if (__all_sync(0xffffffffu, !(is_sentinel(a) || is_sentinel(b)))) break;
            ++spins_total;
            if (spins.expired()) { err = BFD_ETIMEOUT; break; }
            spin_pause();
        }
        if (take) {
#pragma unroll
            for (int r = 0; r < kAhead; ++r)
                if (r < ny) cp_async16(ring0 + (saddr)(r * kRecBytes), io.edge_in + 2 * lane + (size_t)r * kRecDoubles);
        }
    }
    // MAIN: every row this call touches is a grid row for every lane (no range tests): steps kMainLo .. ny - kAhead - 2
    auto halo = [&](auto main_phase) {
        constexpr bool MAIN = decltype(main_phase)::value;
        double na = 0.0, nb = 0.0;
        if (take && (MAIN || (unsigned)(rc + 1) < (unsigned)ny)) lds2(sl, na, nb);
        if (take && (MAIN || (unsigned)rc < (unsigned)ny)) {
            if (is_sentinel_hi(va) || is_sentinel_hi(vb)) {          // had not arrived when the row was copied
                ++misses;
                const PolledPiece p = poll_piece(gin, in_remote);
                va = p.a; vb = p.b; spins_total += p.spins;
                if (!p.ok) err = BFD_ETIMEOUT;
            }
            sts2(x_cur, va, vb);
#ifndef BFD_X_NOREARM
            st_weak2(gin, armed, armed);                             // re-arm for the next launch
#endif
            if (MAIN || rc + kAhead < ny) cp_async16(pl, gin + (size_t)kAhead * kRecDoubles);
        }
        if (!kSplitDuties) { if (MAIN) edge.step_main(); else edge.step(); }
        va = na; vb = nb;
        ++rc; gin += kRecDoubles;
        sl += (saddr)kRecBytes; if (sl == ring_end) sl = ring0;
        pl += (saddr)kRecBytes; if (pl == ring_end) pl = ring0;
        x_cur = x_sum - x_cur;
    };

    // One commit group per step (the feed duty's row rides in it when this warp carries that duty): "all
    // but the newest kPfPending groups" = "the hand-off rows copied >= 3 steps ago have landed".
    if (!kSplitDuties) feed.prologue();
    else { cp_async_commit(); cp_async_wait_pending<0>(); }      // the first kAhead rows of the ring
    if (take && (unsigned)rc < (unsigned)ny) lds2(ring0 + (saddr)((rc & (kInRing - 1)) * kRecBytes), va, vb);
    halo(BoolTag<false>{});                  // "step -1": what lane 0 reads in step 0
    __syncthreads();
    // head, main (no range tests: rows s + 1 - level .. s + kAhead + 1 - level consumed / copied, row s - 33 - level sent,
    // row s + kPfDepth - 1 fed are grid rows for every level 0..K), tail; the two edge phases share one copy of the code
    const int main_lo = kSkew + 2 + K + 1, main_hi = ny - kAhead - 1;
    int s = 0;
#pragma unroll 1
    for (int phase = 0; phase < 2; ++phase) {
        const int ep_end = phase == 0 ? (main_lo < nsteps ? main_lo : nsteps) : nsteps;
#pragma unroll 1
        for (; s < ep_end; ++s) {
            if (!kSplitDuties) feed.step();
            else { cp_async_commit(); cp_async_wait_pending<kPfPending>(); }
            halo(BoolTag<false>{});
            __syncthreads();
        }
        if (phase == 0 && UNI && !kSplitDuties) {
#pragma unroll 2
            for (; s < main_hi; ++s) {
                feed.step_main();
                halo(BoolTag<true>{});
                __syncthreads();
            }
        }
    }
    halo(BoolTag<false>{});                  // what lane 31 published in the last step
    if (io.stats) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            misses += __shfl_xor_sync(0xffffffffu, misses, off);
            spins_total += __shfl_xor_sync(0xffffffffu, spins_total, off);
        }
    }
    if (__any_sync(0xffffffffu, err != 0) && lane == 0) atomicExch(io.error, BFD_ETIMEOUT);
    if (lane == 0 && io.stats) *io.stats = ((unsigned long long)misses << 32) | spins_total;
}

// First level of warp w when the K+1 levels of a pass (K solves + the residual of x_K) are
// split between kWarpsPerStrip = 4 warps; warp w runs levels [part(K, w), part(K, w+1)).
// Balanced by FP64 work: level 0 also forms Au, Av (57 ops), a solve level is 46, the
// residual level 17.
__host__ __device__ constexpr int part(int K, int w) {
    if (kWarpsPerStrip == 1) return w == 0 ? 0 : K + 1;
    if (kWarpsPerStrip == 2) return w == 0 ? 0 : (w == 1 ? (K + 1) / 2 : K + 1);
    if (kWarpsPerStrip == 5) {
        if (K + 1 <= 5) return w < K + 1 ? w : K + 1;
        if (K == 5) return w < 4 ? w : (w == 4 ? 4 : 6);                 // 0 | 1 | 2 | 3 | 4 5
        return w == 0 ? 0 : w == 1 ? 1 : w == 2 ? 2 : w == 3 ? 3 : w == 4 ? 5 : 7;   // K == 6: 0 | 1 | 2 | 3 4 | 5 6
    }
    if (K + 1 <= 4) return w < K + 1 ? w : K + 1;
    if (K == 4) return w == 0 ? 0 : w == 1 ? 1 : w == 2 ? 2 : w == 3 ? 3 : 5;
    if (K == 5) return w == 0 ? 0 : w == 1 ? 1 : w == 2 ? 3 : w == 3 ? 4 : 6;
    return w == 0 ? 0 : w == 1 ? 1 : w == 2 ? 3 : w == 3 ? 5 : 7;     // K == 6
}

template <int K, int WARP, bool CONT, bool UNI>
__device__ __forceinline__ void sweep_warp(const StepIo &io) {
    constexpr int LO = part(K, WARP), HI = part(K, WARP + 1);
    constexpr int DUTY = !kSplitDuties ? kDutyNone : WARP == 1 ? kDutyFeed : WARP == 2 ? kDutyGive : kDutyNone;
    if constexpr (LO < HI) sweep_part<K, LO, HI, CONT, UNI, DUTY>(io);
    else idle_part<K, CONT, UNI, DUTY>(io);
}

// The K+1 levels of a time step split between the warps of the CTA.  DMA: this copy of the code is the one
// the data-movement warp runs (the kernel splits the two kinds of warp at the top: register budgets).
template <int K, bool CONT, bool UNI, bool DMA>
__device__ __forceinline__ void sweep_step(const StepIo &io, int warp) {
    if constexpr (DMA) { dma_part<K, CONT, UNI>(io); return; }
    if (warp == 0) sweep_warp<K, 0, CONT, UNI>(io);
    if constexpr (kWarpsPerStrip > 1) { if (warp == 1) sweep_warp<K, 1, CONT, UNI>(io); }
    if constexpr (kWarpsPerStrip > 2) {
        if (warp == 2) sweep_warp<K, 2, CONT, UNI>(io);
        if (warp == 3) sweep_warp<K, 3, CONT, UNI>(io);
    }
    if constexpr (kWarpsPerStrip > 4) { if (warp == 4) sweep_warp<K, 4, CONT, UNI>(io); }
}

}  // namespace bfd

namespace bfd {

// newton_raphson's stopping rule (hypernet2D.py:1907-1930) applied to the residual norms one
// pass produced: rn[k] = ||r(x_{kb+k})||, k = 0..K.  Run by the last strip to finish the step.
// Returns true when the speculated solve count was wrong and the sample must stop and restart.
__device__ bool validate_step(const FomArgs &a, int m, int i, int kb, int K, double init_prev) {
    const int ns = a.ns_global;
    const volatile double *part = a.partial + (size_t)(m * a.chunk + i) * ns * kNormSlots;
    double rn[kNormSlots];
    for (int k = 0; k <= K; ++k) {
        double sum = 0.0;
        for (int j = 0; j < ns; ++j) sum += part[(size_t)j * kNormSlots + k];
        rn[k] = sqrt(sum);
    }
    StepRec rec;
    rec.init = kb ? init_prev : rn[0];          // init_norm = ||r(x_0)||, x_0 = wp (:1907)
    rec.status = kStepNone; rec.kconv = K; rec.rel = 0.0;
    for (int k = 0; k < K; ++k) {               // the test precedes each solve (:1915-1921)
        const double rel = rn[k] / rec.init;
        if (rel < a.rel_tol) { rec.status = kStepEarly; rec.kconv = k; rec.rel = rel; break; }
    }
    if (rec.status == kStepNone) {
        if (kb + K >= a.max_its) {              // cap: max_its solves applied, no further test
            rec.status = kStepOkCap; rec.rel = rn[K - 1] / rec.init;
        } else {
            rec.rel = rn[K] / rec.init;
            rec.status = rec.rel < a.rel_tol ? kStepOk : kStepNotConverged;
        }
    }
    a.steprec[(size_t)m * a.chunk + i] = rec;
    return rec.status == kStepNotConverged || (rec.status == kStepEarly && a.strict && rec.kconv > 0);
}

__device__ __forceinline__ int clampk(int k, int cap) { return k < 1 ? 1 : (k > cap ? cap : k); }

// Turn the verdicts of a segment (attempts i0 .. i0+n-1 of this launch, time steps t0 ..) into
// the sample's next state.  Run by the last strip of the sample to finish the segment.
__device__ void commit_segment(const FomArgs &a, int m, int t0, int i0, int n, int kbase, int k_first, int k_spec) {
    FomCtl *ctl = a.ctl + m;
    if (n <= 0) return;
    int t = t0, knext = k_spec, redo = 0;
    int new_kbase = 0, new_kfirst = k_first;
    bool stop = false;
    // Speculation policy: a continuation pass (one more sweep in place) is cheaper than redoing a
    // step, so when a step needed one the count is NOT raised at once -- only when two steps in
    // a row needed it.  A solve count that flips between K and K+1 from step to step then costs
    // one continuation pass every other step instead of a redo or a continuation in every step.
    unsigned under = ctl->under;
    auto completed = [&](int kb, int total) {
        if (kb > 0) { ++under; return under >= 2 ? total : kb; }
        under = 0;
        return total;
    };
    for (int ii = 0; ii < n && !stop; ++ii) {
        const volatile StepRec *vr = a.steprec + (size_t)m * a.chunk + i0 + ii;   // written by another CTA: not through L1
        StepRec rec;
        rec.status = vr->status; rec.kconv = vr->kconv; rec.rel = vr->rel; rec.init = vr->init;
        const int kb = ii ? 0 : kbase;
        int K = ii ? k_spec : k_first;
        if (K > a.max_its - kb) K = a.max_its - kb;
        if (K > a.kcap) K = a.kcap;
        int *its = a.its + (size_t)m * a.num_steps + t;
        double *rel = a.relnorm + (size_t)m * a.num_steps + t;
        switch (rec.status) {
        case kStepOk:     *its = kb + K + 1; *rel = rec.rel; knext = completed(kb, kb + K); ++t; break;   // == len(resnorms)
        case kStepOkCap:  *its = a.max_its;  *rel = rec.rel; knext = completed(kb, kb + K); ++t; break;
        case kStepEarly:
            if (!a.strict || rec.kconv == 0) {   // keep the further-converged iterate
                *its = kb + K + 1; *rel = rec.rel; knext = kb + rec.kconv; ++t;
            } else {                              // the reference returns x_kconv: redo with that count
                new_kbase = 0; new_kfirst = rec.kconv; stop = true; ++redo;
                knext = rec.kconv;                // ... and expect it of the steps that follow
            }
            break;
        default:                                  // not converged: further sweeps in place
            new_kbase = kb + K; new_kfirst = 1; ctl->init_norm = rec.init; stop = true; ++redo;
            break;
        }
    }
    if (!stop) { ctl->kbase = 0; ctl->k_first = clampk(knext, a.kcap); ctl->k_spec = clampk(knext, a.kcap); }
    else { ctl->kbase = new_kbase; ctl->k_first = new_kfirst; if (new_kbase == 0) ctl->k_spec = clampk(knext, a.kcap); }
    ctl->t = t;
    ctl->under = under;
    if (redo) { ctl->redone += redo; atomicAdd(&a.glob->redone, redo); }
    if (t >= a.num_steps) {
#ifndef BFD_EMULATE
        if (a.slab) { __threadfence_system(); atomicAdd_system(&a.glob->done, 1); } else
#endif
        atomicAdd(&a.glob->done, 1);
    }
}

// Up to `chunk` time-step attempts of every unfinished sample.  One CTA <-> (sample m, strip j);
// virtual ids come from a ticket so that a strip's left neighbour (a smaller id) is always
// already running when the strip waits on it.
//
// Speculation control.  A sample runs its steps back to back with the speculated solve count.
// The strip that completes a step's norms validates it; on a wrong count it raises the stop
// bit of the sample's gate.  Every strip passes the gate before each attempt (atomicMax with
// the attempt index): attempts some strip had already started are still run by all strips
// (every hand-off record produced is consumed), later ones are refused.  The refused strips
// wait for the commit by the last of them (all strips of the sample are resident by then: they
// all took part in the failed step), which rewinds the sample to the failed step with the
// right count, and carry on inside the same launch.
//
// SLAB: the strips of a sample are spread over several GPUs (bfd_fom_slab).  Control words live
// in rank 0's memory, so their atomics and fences are system-scoped, and since the ranks' launch
// sequences are not ordered by a stream, a launch waits until every strip of every rank has left
// the launch before it (ctl->exits) before it reads the sample's state.
template <bool SLAB> __device__ __forceinline__ void fence() { if (SLAB) __threadfence_system(); else __threadfence(); }
template <bool SLAB> __device__ __forceinline__ unsigned a_inc(unsigned *p, unsigned lim) {
#ifndef BFD_EMULATE
    if (SLAB) return atomicInc_system(p, lim);
#endif
    return atomicInc(p, lim);
}
template <bool SLAB> __device__ __forceinline__ unsigned a_max(unsigned *p, unsigned v) {
#ifndef BFD_EMULATE
    if (SLAB) return atomicMax_system(p, v);
#endif
    return atomicMax(p, v);
}
template <bool SLAB> __device__ __forceinline__ unsigned a_or(unsigned *p, unsigned v) {
#ifndef BFD_EMULATE
    if (SLAB) return atomicOr_system(p, v);
#endif
    return atomicOr(p, v);
}
template <bool SLAB> __device__ __forceinline__ unsigned a_add(unsigned *p, unsigned v) {
#ifndef BFD_EMULATE
    if (SLAB) return atomicAdd_system(p, v);
#endif
    return atomicAdd(p, v);
}

constexpr int kChunkSmemDoubles = kInRing * kRecDoubles + (kHandIfaces * kHandBytes + kXchgBytes + kPfBytes) / 8;

// The work of one CTA of fom_chunk_kernel, as the compute warps (DMA = false) or the data-movement warp
// (DMA = true) run it: the same control flow and the same barriers, two copies of the code (the commit logic
// of the leader thread exists in the compute copy only, the record hand-off in the other).
template <bool UNI, bool SLAB, bool DMA>
__device__ __forceinline__ void chunk_main(const FomArgs &a, double *s_mem, int *s_ctl, double *s_dbl, int item,
                                           const int warp) {
    const int lane = lane_id();
    const bool leader = !DMA && warp == 0 && lane == 0;     // `warp` is the role: 0 .. kWarpsPerStrip-1 compute, then data movement
    const int nstrips = a.ssl.nstrips;
    const int m = item / nstrips, j = item - m * nstrips;
    const int jg = j + a.strip0, nsg = a.ns_global;        // this strip and the strip count in the whole grid
    FomCtl *ctl = a.ctl + m;
    if (SLAB) {
        if (leader && a.launch_index > 0) {
            const unsigned need = (unsigned)a.launch_index * (unsigned)nsg;
            SpinClock spins;
            while (*(volatile unsigned *)&ctl->exits < need && *(volatile int *)&ctl->t < a.num_steps &&
                   *(volatile int *)&a.glob->error == 0) {
                spin_pause();
                if (spins.expired()) { atomicExch(&a.glob->error, BFD_ETIMEOUT); break; }
            }
            fence<SLAB>();
        }
        __syncthreads();
    }

    const int nx = a.ssl.nx, ny = a.ssl.ny;
    const size_t sd = a.ssl.state_doubles;
    const size_t strip_off = (size_t)j * a.ssl.strip_stride + (size_t)lane;
    double *pool_m = a.pool + (size_t)m * a.pool_stride_mu + strip_off;
    double *edge_m = a.edge + (size_t)m * a.edge_stride_mu;

    StepIo io;
    io.ring_in = smem_addr(s_mem);
    const saddr hand0 = io.ring_in + (saddr)(kInRing * kRecBytes);
    io.hand = hand0 + (saddr)(warp * kHandBytes);
    io.xchg = hand0 + (saddr)(kHandIfaces * kHandBytes);
    io.pf = io.xchg + (saddr)kXchgBytes;
    io.lbc = a.lbc + (size_t)m * ny;
    io.cy = a.cy; io.cyd = a.cyd;
    io.ny = ny; io.strip = jg; io.nstrips = nsg; io.lane = lane;
    io.in_remote = SLAB && j == 0 && jg > 0;
    io.out_remote = SLAB && j == nstrips - 1 && jg < nsg - 1;
    const int c = j * 32 + lane;
    io.colok = c < nx;
    const int cc = io.colok ? c : nx - 1;
    io.cx = a.cx[cc]; io.cxl = a.cxl[cc]; io.src = a.src[(size_t)m * nx + cc];
    io.inlet = (jg == 0 && lane == 0);
    io.error = &a.glob->error;

    int attempt = 0;                 // attempts of this launch so far (same on every strip of the sample)
    for (;;) {                       // segments: runs of steps between two commits
        __syncthreads();
        unsigned epoch0 = 0;
        if (leader) {
            // control words are read by one thread and broadcast by shuffle: the compiler then
            // knows they are warp-uniform
            const volatile FomCtl *vc = ctl;       // written by other CTAs (other GPUs in slab runs)
            s_ctl[1] = vc->t; s_ctl[2] = vc->kbase; s_ctl[3] = vc->k_first; s_ctl[4] = vc->k_spec;
            epoch0 = vc->epoch;
            s_dbl[0] = vc->init_norm;
        }
        __syncthreads();
        const int w = (lane < 8) ? s_ctl[lane] : 0;
        const int t0 = __shfl_sync(0xffffffffu, w, 1);
        if (t0 >= a.num_steps || attempt >= a.chunk) break;   // trajectory finished / budget of this launch used
        const int kbase = __shfl_sync(0xffffffffu, w, 2), k_first = __shfl_sync(0xffffffffu, w, 3),
                  k_spec = __shfl_sync(0xffffffffu, w, 4);
        const double init_prev = s_dbl[0];
        const int i0 = attempt;
        bool refused = false;
        for (; attempt < a.chunk; ++attempt) {
            const int ii = attempt - i0, t = t0 + ii;
            if (t >= a.num_steps) break;
            __syncthreads();
            if (leader) {
                const unsigned old = a_max<SLAB>(&ctl->gate, (unsigned)attempt);
                s_ctl[5] = ((old & kGateStop) && (unsigned)attempt > (old & (kGateStop - 1))) ? 0 : 1;
            }
            __syncthreads();
            const int gw = (lane < 8) ? s_ctl[lane] : 0;
            if (__shfl_sync(0xffffffffu, gw, 5) == 0) { refused = true; break; }

            const int i = attempt;
            const int kb = ii ? 0 : kbase;
            int K = ii ? k_spec : k_first;
            if (K > a.max_its - kb) K = a.max_its - kb;
            io.xwp = pool_m + (size_t)t * sd;
            io.xout = pool_m + (size_t)(t + 1) * sd;
            io.xin = kb ? io.xout : io.xwp;
            const size_t eoff = (size_t)i * ny * kRecDoubles;
            io.edge_in = edge_m + (size_t)(j > 0 ? j - 1 : 0) * a.edge_stride_strip + eoff;
            io.edge_out = edge_m + (size_t)(j < nstrips - 1 ? j : 0) * a.edge_stride_strip + eoff;
            if (SLAB) {   // the records that cross a slab boundary live in the consumer's halo buffer
                if (io.in_remote) io.edge_in = a.halo + (size_t)m * a.edge_stride_strip + eoff;
                if (io.out_remote) io.edge_out = a.next_halo + (size_t)m * a.edge_stride_strip + eoff;
            }
            io.partial = a.partial + ((size_t)(m * a.chunk + i) * nsg + jg) * kNormSlots;
            unsigned long long *tr = a.trace + ((size_t)(m * a.chunk + i) * nstrips + j) * 3;
            if (leader) tr[0] = global_timer_ns();
            io.stats = tr + 2;
            if (kb) {
                sweep_step<1, true, UNI, DMA>(io, warp);
            } else {
                switch (K) {
                case 1: sweep_step<1, false, UNI, DMA>(io, warp); break;
                case 2: sweep_step<2, false, UNI, DMA>(io, warp); break;
                case 3: sweep_step<3, false, UNI, DMA>(io, warp); break;
                case 4: sweep_step<4, false, UNI, DMA>(io, warp); break;
                case 5: sweep_step<5, false, UNI, DMA>(io, warp); break;
                default: sweep_step<6, false, UNI, DMA>(io, warp); break;
                }
            }
            __syncthreads();       // every warp has published its norms
            if (leader) {
                tr[1] = global_timer_ns();
                fence<SLAB>();
                if (a_inc<SLAB>(&a.arrive[(size_t)m * a.chunk + i], (unsigned)nsg - 1) == (unsigned)nsg - 1) {
                    fence<SLAB>();
                    if (validate_step(a, m, i, kb, kb ? 1 : K, init_prev)) a_or<SLAB>(&ctl->gate, kGateStop);
                }
            }
        }
        // end of the segment: the last strip commits; refused strips wait for it and restart
        if (leader) {
            fence<SLAB>();
            if (a_inc<SLAB>(&ctl->arrive, (unsigned)nsg - 1) == (unsigned)nsg - 1) {
                fence<SLAB>();
                commit_segment(a, m, t0, i0, attempt - i0, kbase, k_first, k_spec);
                // next segment of this launch continues at `attempt`; a new launch starts at 0
                *(volatile unsigned *)&ctl->gate = refused ? (unsigned)(attempt > 0 ? attempt - 1 : 0) : 0u;
                fence<SLAB>();
                a_add<SLAB>(&ctl->epoch, 1u);
            } else if (refused) {
                SpinClock spins;
                while (*(volatile unsigned *)&ctl->epoch == epoch0) {
                    spin_pause();
                    if (spins.expired()) { atomicExch(&a.glob->error, BFD_ETIMEOUT); break; }
                }
                fence<SLAB>();
            }
        }
        if (!refused) break;
    }
    if (SLAB) {   // this strip is done with the launch: its re-armed records and commits are visible
        __syncthreads();
        if (leader) { fence<SLAB>(); a_add<SLAB>(&ctl->exits, 1u); }
    }
}

template <bool UNI, bool SLAB>
__global__ void __launch_bounds__(32 * (kWarpsPerStrip + 1), BFD_MINB) fom_chunk_kernel(const FomArgs a) {
    BFD_SHARED_ARRAY(double, s_mem, [kChunkSmemDoubles]);
    BFD_SHARED_ARRAY(int, s_ctl, [8]);
    BFD_SHARED_ARRAY(double, s_dbl, [2]);
    for (int i = threadIdx.x; i < kChunkSmemDoubles; i += blockDim.x) s_mem[i] = 0.0;
    if (threadIdx.x == 0) s_ctl[0] = (int)atomicInc(a.ticket, gridDim.x - 1);
    __syncthreads();
    const int item = __shfl_sync(0xffffffffu, s_ctl[0], 0);
    if (item >= a.batch * a.ssl.nstrips) return;
    // Roles by warp index.  The 5 warps of a CTA sit on consecutive hardware warp slots and slot mod 4 is the SM
    // sub-partition, so with 4 CTAs per SM every sub-partition holds one warp of each role; giving every
    // sub-partition ONE compute role instead (roles by %warpid: two loop bodies in its instruction cache instead
    // of five) was measured 19 % slower -- the heavy roles then share an FP64 pipe (profiles/r02_notes.md, 8).
    const int role = threadIdx.x >> 5;
    if (role == kWarpsPerStrip) chunk_main<UNI, SLAB, true>(a, s_mem, s_ctl, s_dbl, item, role);
    else chunk_main<UNI, SLAB, false>(a, s_mem, s_ctl, s_dbl, item, role);
}

}  // namespace bfd

#include "fom_tile.cuh"

namespace bfd {

// natural (u|v, row-major) -> skewed strip layout, into every sample's slot 0
__global__ void fom_import_kernel(const FomArgs a, const double *w0, size_t w0_stride) {
    const Ssl &L = a.ssl;
    const size_t per_state = L.state_doubles;
    const size_t total = per_state * a.batch;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total;
         i += (size_t)gridDim.x * blockDim.x) {
        const int m = (int)(i / per_state);
        size_t o = i - (size_t)m * per_state;
        const size_t o_state = o;
        const int strip = (int)(o / L.strip_stride);
        o -= (size_t)strip * L.strip_stride;
        const int s = (int)(o / (2 * L.W));
        const int rem = (int)(o - (size_t)s * 2 * L.W);
        const int comp = rem / L.W, cc = rem % L.W;
        const int r = s - cc / L.cpl, c = strip * L.W + cc;
        double val = 0.0;
        if (r >= 0 && r < L.ny && c < L.nx)
            val = w0[(size_t)m * w0_stride + (size_t)comp * L.nx * L.ny + (size_t)r * L.nx + c];
        a.pool[(size_t)m * a.pool_stride_mu + o_state] = val;
    }
}

// arm the hand-off records and reset the per-sample control blocks
// `control` = 0 on the ranks of a slab run that do not own the control blocks
__global__ void fom_reset_kernel(const FomArgs a, int control) {
    const size_t n_edge = a.edge_stride_mu * a.batch;
    const double sentinel = __longlong_as_double((long long)kSentinelBits);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n_edge;
         i += (size_t)gridDim.x * blockDim.x)
        a.edge[i] = sentinel;
    const size_t n_halo = a.halo ? a.edge_stride_strip * a.batch : 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n_halo;
         i += (size_t)gridDim.x * blockDim.x)
        a.halo[i] = sentinel;
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g == 0) *a.ticket = 0;
    if (!control) return;
    const size_t n_arrive = (size_t)a.batch * a.chunk;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n_arrive;
         i += (size_t)gridDim.x * blockDim.x)
        a.arrive[i] = 0u;
    if (g == 0) {
        a.glob->done = (a.num_steps == 0) ? a.batch : 0; a.glob->error = 0;
        a.glob->redone = 0;
    }
    if (g < a.batch) {
        FomCtl c;
        c.t = 0; c.kbase = 0; c.k_first = a.hint < a.kcap ? a.hint : a.kcap; c.k_spec = c.k_first;
        c.gate = 0; c.arrive = 0; c.redone = 0; c.epoch = 0; c.init_norm = 0.0;
        c.exits = 0; c.under = 0; c.pad1[0] = c.pad1[1] = 0.0;
        a.ctl[g] = c;
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
    const double *pool_m = a.pool + (size_t)m * a.pool_stride_mu;
    for (int comp = 0; comp < 2; ++comp) {
        for (int cc0 = 0; cc0 < L.W; cc0 += 32) {
            for (int tb = 0; tb < nt; tb += 32) {
                // load: tile[time][entry]
                for (int ty = threadIdx.y; ty < 32; ty += blockDim.y) {
                    const int t = tb + ty;
                    if (t < nt) {
                        const size_t o = (size_t)(t0 + t) * L.state_doubles +
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

static int fom_init_impl(const bfd_fom_desc *d, const SlabView &v, void *ws, size_t ws_bytes, void *stream) {
    if (!ws) return BFD_EINVAL;
    const FomLayout L = fom_layout(d, v);
    if (ws_bytes < L.total) return BFD_EWORKSPACE;
    const FomArgs a = fom_args(d, L, ws, v);
    BFD_LAUNCH(fom_reset_kernel, BFD_GRID(592), 256, 0, (cudaStream_t)stream, a, (!v.slab || v.rank == 0) ? 1 : 0);
    BFD_LAUNCHED();
    return BFD_OK;
}

static int fom_set_initial_impl(const bfd_fom_desc *d, const SlabView &v, void *ws, const double *w0,
                                size_t w0_stride, void *stream) {
    if (!ws || !w0) return BFD_EINVAL;
    const FomLayout L = fom_layout(d, v);
    const FomArgs a = fom_args(d, L, ws, v);
    BFD_LAUNCH(fom_import_kernel, BFD_GRID(592), 256, 0, (cudaStream_t)stream, a, w0, w0_stride);
    BFD_LAUNCHED();
    return BFD_OK;
}

// The run's status words, written by the GPU straight into the caller's page-locked memory (zero-copy):
// polling must not go through a copy engine, where an 8-byte cudaMemcpyAsync queues behind whatever bulk
// device-to-host copies the process has in flight -- a sweep that streams 4.5 GB trajectories to the host held
// every poll back by one or more 79 ms copies and ran its launches one at a time (profiles/r02_notes.md, 9).
__global__ void fom_status_kernel(const FomGlobal *g, int *host_words) {
    const volatile FomGlobal *v = g;
    host_words[0] = v->done;
    host_words[1] = v->error;
    host_words[2] = v->redone;
    __threadfence_system();
}

static int launch_chunk(const FomArgs &a, cudaStream_t st) {
    const int grid = a.batch * a.ssl.nstrips;
    if (a.variant == 1) {
        void (*kernel)(const FomArgs) =
            a.slab ? (a.uniform_y ? fom_tile_kernel<true, true> : fom_tile_kernel<false, true>)
                   : (a.uniform_y ? fom_tile_kernel<true, false> : fom_tile_kernel<false, false>);
        BFD_LAUNCH_CORESIDENT(kernel, grid, 32 * kTileDepth, 0, st, a);
        BFD_LAUNCHED();
        return BFD_OK;
    }
    const int block = 32 * (kWarpsPerStrip + 1);
    void (*kernel)(const FomArgs) =
        a.slab ? (a.uniform_y ? fom_chunk_kernel<true, true> : fom_chunk_kernel<false, true>)
               : (a.uniform_y ? fom_chunk_kernel<true, false> : fom_chunk_kernel<false, false>);
    BFD_LAUNCH_CORESIDENT(kernel, grid, block, 0, st, a);
    BFD_LAUNCHED();
    return BFD_OK;
}

static int fom_run_impl(const bfd_fom_desc *d, const SlabView &v, void *ws, const bfd_fom_coef *coef, void *h_pinned,
                        void *stream) {
    if (!ws || !coef || !h_pinned) return BFD_EINVAL;
    if (v.slab && ((v.rank > 0 && !v.home) || (v.rank < v.world - 1 && !v.next_halo))) return BFD_EINVAL;
    const FomLayout L = fom_layout(d, v);
    FomArgs a = fom_args(d, L, ws, v);
    a.cx = coef->cx; a.cxl = coef->cxl; a.cy = coef->cy; a.cyd = coef->cyd;
    a.src = coef->src; a.lbc = coef->lbc;
    cudaStream_t st = (cudaStream_t)stream;
    volatile int *h = (volatile int *)h_pinned;
    h[0] = 0; h[1] = 0; h[2] = 0; h[3] = 0;
    if (d->num_steps == 0) { h[0] = d->batch; return BFD_OK; }

    // Samples advance independently on the device; the host keeps two launches queued and looks
    // at the "samples finished" counter of the launch before them, so the GPU never idles on
    // the poll.  A launch issued after everything finished exits at once.  (Slab runs: the
    // counter is rank 0's; ranks may issue different numbers of trailing empty launches.)
    cudaEvent_t ev[2];
    BFD_CUDA(cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming));
    BFD_CUDA(cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming));
    int *h_poll = (int *)h_pinned + 4;   // two landing zones of 3 ints at h[4..6], h[7..9]; final status at h[10..12]
    int *d_poll = nullptr;               // the same words as the device sees them; null: not mapped, copy instead
    if (cudaHostGetDevicePointer((void **)&d_poll, h_poll, 0) != cudaSuccess) { cudaGetLastError(); d_poll = nullptr; }
    const long long hard_cap = (long long)d->num_steps * d->max_its + 8;
    long long launched = 0;
    int rc = BFD_OK, pending = 0, slot = 0;
    bool finished = false;
    while (!finished && rc == BFD_OK) {
        a.launch_index = (int)launched;
        rc = launch_chunk(a, st);
        ++launched;
        if (rc != BFD_OK) break;
        cudaError_t e = cudaSuccess;
        if (d_poll) {
            BFD_LAUNCH(fom_status_kernel, 1, 1, 0, st, a.glob, d_poll + 3 * slot);
            BFD_LAUNCHED();
        } else {
            e = cudaMemcpyAsync(h_poll + 3 * slot, &a.glob->done, 3 * sizeof(int), cudaMemcpyDeviceToHost, st);
        }
        if (e == cudaSuccess) e = cudaEventRecord(ev[slot], st);
        if (e != cudaSuccess) { rc = record_cuda_error(e, "fom poll"); break; }
        if (pending) {   // look at the poll issued one launch ago
            const int prev = slot ^ 1;
            e = cudaEventSynchronize(ev[prev]);
            if (e != cudaSuccess) { rc = record_cuda_error(e, "fom poll sync"); break; }
            const volatile int *hp = h_poll + 3 * prev;
            if (hp[0] >= d->batch || hp[1] != 0) finished = true;
        }
        pending = 1;
        slot ^= 1;
        if (launched > hard_cap) { rc = BFD_ENOTCONVERGED; break; }
    }
    if (rc == BFD_OK && d_poll) {
        BFD_LAUNCH(fom_status_kernel, 1, 1, 0, st, a.glob, d_poll + 6);
        BFD_LAUNCHED();
    }
    cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess && rc == BFD_OK) rc = record_cuda_error(e, "fom run sync");
    cudaEventDestroy(ev[0]);
    cudaEventDestroy(ev[1]);
    if (rc == BFD_OK) {
        int fin[4] = {0, 0, 0, 0};
        if (d_poll) {
            const volatile int *hp = h_poll + 6;
            fin[0] = hp[0]; fin[1] = hp[1]; fin[2] = hp[2];
        } else {
            e = cudaMemcpy(fin, &a.glob->done, 3 * sizeof(int), cudaMemcpyDeviceToHost);
            if (e != cudaSuccess) return record_cuda_error(e, "fom final status");
        }
        h[0] = fin[0]; h[1] = fin[1]; h[2] = (int)launched; h[3] = fin[2];
        if (fin[1] != 0) rc = fin[1];
        else if (fin[0] < d->batch) rc = BFD_ENOTCONVERGED;
    }
    return rc;
}

static int fom_export_impl(const bfd_fom_desc *d, const SlabView &v, const void *ws, int mu_index, int t0, int t1,
                           int layout, size_t ld, double *out, void *stream) {
    if (!ws || !out) return BFD_EINVAL;
    if (mu_index < 0 || mu_index >= d->batch || t0 < 0 || t1 > d->num_steps + 1 || t0 >= t1) return BFD_EINVAL;
    if (layout != 0 && layout != 1) return BFD_EINVAL;
    if (layout == 1 && ld < (size_t)(t1 - t0)) return BFD_EINVAL;
    const FomLayout L = fom_layout(d, v);
    const FomArgs a = fom_args(d, L, const_cast<void *>(ws), v);
    dim3 grid(d->ny + kSkew, L.ssl.nstrips), block(32, 8);
    BFD_LAUNCH(fom_export_kernel, grid, block, 0, (cudaStream_t)stream, a, mu_index, t0, t1, layout, ld, out);
    BFD_LAUNCHED();
    return BFD_OK;
}

static int fom_diagnostics_impl(const bfd_fom_desc *d, const SlabView &v, const void *ws, int32_t *its,
                                double *relnorm, void *stream) {
    if (!ws) return BFD_EINVAL;
    const FomLayout L = fom_layout(d, v);
    const FomArgs a = fom_args(d, L, const_cast<void *>(ws), v);
    const size_t n = (size_t)d->batch * d->num_steps;
    if (its) BFD_CUDA(cudaMemcpyAsync(its, a.its, n * sizeof(int), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    if (relnorm)
        BFD_CUDA(cudaMemcpyAsync(relnorm, a.relnorm, n * sizeof(double), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return BFD_OK;
}

// ---- whole grid on one GPU ----------------------------------------------------------------
extern "C" size_t bfd_fom_workspace_bytes(const bfd_fom_desc *d) {
    if (check_desc(d) != BFD_OK) return 0;
    return fom_layout(d, whole_grid(d)).total;
}

extern "C" int bfd_fom_init(const bfd_fom_desc *d, void *ws, size_t ws_bytes, void *stream) {
    if (check_desc(d) != BFD_OK) return BFD_EINVAL;
    return fom_init_impl(d, whole_grid(d), ws, ws_bytes, stream);
}

extern "C" int bfd_fom_set_initial(const bfd_fom_desc *d, void *ws, const double *w0, size_t w0_stride,
                                   void *stream) {
    if (check_desc(d) != BFD_OK) return BFD_EINVAL;
    return fom_set_initial_impl(d, whole_grid(d), ws, w0, w0_stride, stream);
}

extern "C" int bfd_fom_run(const bfd_fom_desc *d, void *ws, const bfd_fom_coef *coef, void *h_pinned,
                           void *stream) {
    if (check_desc(d) != BFD_OK) return BFD_EINVAL;
    return fom_run_impl(d, whole_grid(d), ws, coef, h_pinned, stream);
}

extern "C" int bfd_fom_export(const bfd_fom_desc *d, const void *ws, int mu_index, int t0, int t1, int layout,
                              size_t ld, double *out, void *stream) {
    if (check_desc(d) != BFD_OK) return BFD_EINVAL;
    return fom_export_impl(d, whole_grid(d), ws, mu_index, t0, t1, layout, ld, out, stream);
}

extern "C" int bfd_fom_diagnostics(const bfd_fom_desc *d, const void *ws, int32_t *its, double *relnorm,
                                   void *stream) {
    if (check_desc(d) != BFD_OK) return BFD_EINVAL;
    return fom_diagnostics_impl(d, whole_grid(d), ws, its, relnorm, stream);
}

extern "C" int bfd_fom_trace(const bfd_fom_desc *d, const void *ws, uint64_t *out, void *stream) {
    if (check_desc(d) != BFD_OK || !ws || !out) return BFD_EINVAL;
    const SlabView v = whole_grid(d);
    const FomLayout L = fom_layout(d, v);
    const FomArgs a = fom_args(d, L, const_cast<void *>(ws), v);
    const size_t n = (size_t)d->batch * L.chunk * L.ssl.nstrips * 3;
    BFD_CUDA(cudaMemcpyAsync(out, a.trace, n * sizeof(uint64_t), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return BFD_OK;
}

// ---- one grid over several GPUs: this rank's slab ------------------------------------------
extern "C" size_t bfd_fom_slab_workspace_bytes(const bfd_fom_desc *d, const bfd_fom_slab *sl) {
    SlabView v;
    if (slab_view(d, sl, &v) != BFD_OK) return 0;
    return fom_layout(d, v).total;
}

extern "C" size_t bfd_fom_slab_halo_offset(const bfd_fom_desc *d, const bfd_fom_slab *sl) {
    SlabView v;
    if (slab_view(d, sl, &v) != BFD_OK || v.strip0 == 0) return 0;
    return fom_layout(d, v).off_halo;
}

extern "C" int bfd_fom_slab_init(const bfd_fom_desc *d, const bfd_fom_slab *sl, void *ws, size_t ws_bytes,
                                 void *stream) {
    SlabView v;
    if (slab_view(d, sl, &v) != BFD_OK) return BFD_EINVAL;
    v.home = nullptr;          // a rank resets what it owns: records and ticket; rank 0 also the control blocks
    return fom_init_impl(d, v, ws, ws_bytes, stream);
}

extern "C" int bfd_fom_slab_set_initial(const bfd_fom_desc *d, const bfd_fom_slab *sl, void *ws, const double *w0,
                                        size_t w0_stride, void *stream) {
    SlabView v;
    if (slab_view(d, sl, &v) != BFD_OK) return BFD_EINVAL;
    return fom_set_initial_impl(d, v, ws, w0, w0_stride, stream);
}

extern "C" int bfd_fom_slab_run(const bfd_fom_desc *d, const bfd_fom_slab *sl, void *ws, const bfd_fom_coef *coef,
                                void *h_pinned, void *stream) {
    SlabView v;
    if (slab_view(d, sl, &v) != BFD_OK) return BFD_EINVAL;
    return fom_run_impl(d, v, ws, coef, h_pinned, stream);
}

extern "C" int bfd_fom_slab_export(const bfd_fom_desc *d, const bfd_fom_slab *sl, const void *ws, int mu_index,
                                   int t0, int t1, int layout, size_t ld, double *out, void *stream) {
    SlabView v;
    if (slab_view(d, sl, &v) != BFD_OK) return BFD_EINVAL;
    return fom_export_impl(d, v, ws, mu_index, t0, t1, layout, ld, out, stream);
}

extern "C" int bfd_fom_slab_diagnostics(const bfd_fom_desc *d, const bfd_fom_slab *sl, const void *ws, int32_t *its,
                                        double *relnorm, void *stream) {
    SlabView v;
    if (slab_view(d, sl, &v) != BFD_OK) return BFD_EINVAL;
    return fom_diagnostics_impl(d, v, ws, its, relnorm, stream);
}
