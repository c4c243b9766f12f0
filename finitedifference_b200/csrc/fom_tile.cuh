// fom_tile_kernel: the FOM sweep with one WARP per (strip, time step) and kTileDepth consecutive time
// steps of a strip in flight inside one CTA.  Included by fom.cu (shares FomArgs, the verdict /
// commit logic and the strip-to-strip hand-off records with fom_chunk_kernel).  Selectable variant
// (bfd_fom_desc.kernel = 1); fom_chunk_kernel stays the default because it is faster on B200
// (37.2, later 39.4, vs 34.5-35.2 G cell-updates/s at the bench's shape) -- what was tried and measured is in
// profiles/r02_notes.md and DESIGN.md 4.1b.
//
// Why it was built: fom_chunk_kernel splits the K+1 Newton levels of a time step between four warps
// and a data-movement warp that meet at one CTA barrier per wavefront step.  ncu: every scheduler
// holds 5 warps (one of each role: the warp slots of a 5-warp CTA are dealt round-robin to the four
// schedulers, tools/warp_slot_probe.cu) that each issue once per ~10.6 cycles (barrier 2.6, fixed-
// latency waits 2.4, shared-memory scoreboard 1.1 per issue), FP64 pipe 51 % busy; the register file
// (96 x 160 x 4) admits no more warps.  A variant with ALL levels in one warp reached 64 % of that
// kernel's throughput with a single compute warp per scheduler.  So: give every warp all levels of a
// time step (no level hand-off, no CTA barrier, no data-movement warp) and get the warps from TIME
// instead of from levels -- warp d of a CTA advances time steps t = d, d + D, d + 2D, ... of its strip,
// each trailing its predecessor by a few rows.  The iterate x_{t+1} a warp produces goes to the
// snapshot pool; the warp that advances step t+1 fetches it from there a few rows later as soon as
// the producer's row counter allows.
// What the measurements said: the warps do issue twice as often (5.4 cycles per instruction), but a
// strip-step costs 480-500 instructions instead of 405 (the hand-off and prefetch code of the data-
// movement warp now runs in every warp, ~120; 55-60 register moves for the level pipeline at the 168-
// register cap) and ptxas emits the five independent levels one after the other (fixed-latency waits
// 1.7-2.2 per issue), so three warps per scheduler keep the FP64 pipe 47-48 % busy -- no better than
// five single-level warps with a barrier.
//
//   * level-0 input: a ring of kTRing slot-rows (u | v, 512 B each) per warp, filled kTPf - 1 rows
//     ahead by 16-byte cp.async copies; rows of a time step produced in this segment are gated by the
//     producing warp's row counter in shared memory (monotone stream positions, polled only when the
//     cached value is not enough).  (A first version handed the rows from warp to warp through
//     shared-memory rings: the wrap-around link, last warp -> first warp's NEXT time step, would have to
//     buffer a whole time step and the cyclic chain deadlocks.)
//   * left neighbour inside the strip: the double-buffered exchange rows of fom_chunk_kernel (slot l+1
//     written by lane l, slot l read by lane l one step later), one __syncwarp per step;
//   * strip-to-strip hand-off: the same self-validating 16-byte records in global memory, moved by
//     lanes < NPR of the same warp (piece p by lane p): copied into a shared-memory ring kTAhead rows
//     early, dropped into the halo slot 0 one step early, lane 31's slot 32 sent on the step after;
//   * Newton bookkeeping (speculated solve count, device-side verdict, stop / restart inside the
//     launch) unchanged; at most kTileKMax solves are fused per pass, a step that needs more gets
//     continuation passes.
#pragma once

namespace bfd {

constexpr int kTileDepth = 3;                 // warps per CTA = time steps of one strip in flight
constexpr int kTileKMax = 4;                  // Newton solves fused into one pass
constexpr int kTRing = 8;                     // slot-rows in a warp's level-0 input ring (power of two)
constexpr int kTEdgeRing = 16;                // hand-off record rows staged in shared memory (power of two)
constexpr int kTPieces = 2 * kTileKMax + 1;   // exchange rows = 16-byte pieces per record row (CONT: 4 with K = 1)
constexpr int kTPf = 6;                       // rows of level-0 input in flight when it comes from HBM
constexpr int kTXchgBytes = kTPieces * 2 * kXRow;        // [piece][2 buffers][33 slots x 16 B]
constexpr int kTRingBytes = kTRing * 512;                // [row][u | v][32 lanes x 8 B]
constexpr int kTEdgeRowBytes = kTPieces * 16;
constexpr int kTEdgeBytes = kTEdgeRing * kTEdgeRowBytes;
constexpr int kTWarpBytes = kTXchgBytes + kTRingBytes + kTEdgeBytes;
constexpr int kTWpOffset = 4 * 2 * kXRow;               // CONT (K = 1) keeps the previous snapshot's ring in the unused exchange rows
static_assert(kTWpOffset + kTRingBytes <= kTXchgBytes, "the wp ring of continuation passes fits the spare exchange rows");
static_assert(kTileDepth * kTWarpBytes + 256 <= 48 * 1024, "static shared memory of a CTA");
constexpr int kTAhead = 8;                    // rows copied into the hand-off ring ahead of their use
constexpr int kTStartRow = 24;                // a strip starts a time step when its left neighbour has finished this row
constexpr int kTPublish = 4;                  // the row counter is published every kTPublish rows
static_assert(kTAhead + 2 <= kTEdgeRing, "rows in flight plus the row in use must fit the ring");

struct TileIo {
    const double *xin;    // level-0 iterate when read from HBM: snapshot t, or the partly converged slot t+1 (CONT)
    const double *xwp;    // snapshot t (CONT)
    double *xout;         // slot t+1 (all three offset to this strip and lane)
    double *edge_in, *edge_out;
    saddr xchg, ring, edge_ring;                  // this warp's shared memory
    volatile int *prod_in, *prod_out;             // row counters (stream positions, shared memory): the producer of this step's input; this warp's own
    int in_base, out_base;                        // stream position of slot-row 0: of the input's producer, of this time step
    bool gated, publish;                          // the input is being produced in this segment / somebody will read this step's rows
    double *partial;
    const double *lbc, *cy, *cyd;
    int ny, strip, nstrips, lane;
    double cx, cxl, src;
    bool colok, inlet, in_remote, out_remote;
    int *error;
    unsigned long long *wait_ns;   // trace: time spent waiting for the left strip's start distance
};

#ifdef BFD_EMULATE
#define BFD_TRACE_SPIN(spins, ...) do { if ((spins) == 20000u && emu::tls.lane == 0) { fprintf(stderr, __VA_ARGS__); } } while (0)
#else
#define BFD_TRACE_SPIN(spins, ...) do { } while (0)
#endif
#ifndef BFD_EMULATE
__device__ __forceinline__ void start_pause() { __nanosleep(400); }
__device__ __forceinline__ int ld_shared_volatile(const volatile int *p) { return *p; }
#else
inline void start_pause() { std::this_thread::yield(); }
inline int ld_shared_volatile(const volatile int *p) { std::this_thread::yield(); return __atomic_load_n((const int *)p, __ATOMIC_ACQUIRE); }
#endif

// A value the compiler must keep in a register instead of recomputing it where it is used (under
// register pressure it rematerialises shared-window addresses with S2R SR_CgaCtaId + arithmetic in
// every wavefront step: profiles/r02_notes.md).
// keep_reg: wavefront.cuh

// All K+1 levels of one trapezoidal time step of one strip, by one warp.  Arithmetic identical to
// sweep_part up to exact power-of-two rescalings (the 0.5 factors ride on the iterate): the same bits.
template <int K, bool CONT, bool UNI>
__device__ __forceinline__ void tile_step(const TileIo &io) {
    constexpr int NPR = 2 * K + 1 + (CONT ? 1 : 0);
    constexpr int W = 32;
    static_assert(NPR <= kTPieces && K <= kTileKMax, "exchange rows");
    const int lane = keep_reg(io.lane), ny = io.ny;
    const int last_row = ny + kSkew - 1;   // last slot-row holding cells
    const int nsteps = ny + kSkew + K;
    const bool colok = io.colok;
    const bool gated = io.gated, publish = io.publish;
    const saddr io_xchg = keep_reg(io.xchg), io_ring = keep_reg(io.ring), io_edge_ring = keep_reg(io.edge_ring);

    // 0.5 factors ride on the iterate (hu = u/2, hv = v/2, hpm = (v du + u dv)/2) and signs on the FMA
    // operands: five column / row constants in registers instead of eleven, the same bits
    const double cx = io.cx, cxl = io.cxl;
    const double src = io.src;
    double cyU = 0.0, cydU = 0.0;
    if (UNI) { cyU = __ldg(io.cy + (ny > 1 ? 1 : 0)); cydU = __ldg(io.cyd + (ny > 1 ? 1 : 0)); }

    Level prev[K + 1];
    double qGp = 0.0, qHp = 0.0;
#pragma unroll
    for (int k = 0; k <= K; ++k) prev[k] = Level{0.0, 0.0, 0.0, 0.0};
    double xu[K + 1], xv[K + 1], Au[K + 1], Av[K + 1], Cy[K + 1], Cyd[K + 1], nrm[K + 1];
#pragma unroll
    for (int k = 0; k <= K; ++k) { xu[k] = xv[k] = Au[k] = Av[k] = 0.0; Cy[k] = cyU; Cyd[k] = cydU; nrm[k] = 0.0; }
    const saddr xc_own = io_xchg + (saddr)((lane + 1) * 16);
    const saddr xc_left = io_xchg + (saddr)(lane * 16);
    auto xc_piece = [](int p, int buf) { return (saddr)((p * 2 + buf) * kXRow); };

    // ---- strip-to-strip hand-off: lane p moves piece p of the row records (see dma_part) --------
    const bool special = CONT && lane == 2 * K + 1;
    const int lvl = special ? 0 : (lane >> 1);
    const bool take = lane < NPR && io.strip > 0, give = lane < NPR && io.strip < io.nstrips - 1;
    const bool in_remote = io.in_remote, out_remote = io.out_remote;
    const saddr px = io_xchg + (saddr)((lane < NPR ? lane : 0) * 2 * kXRow);     // piece `lane`, buffer 0
    const saddr ering = io_edge_ring + (saddr)((lane < NPR ? lane : 0) * 16);
    auto ering_row = [&](int r) { return ering + (saddr)((r & (kTEdgeRing - 1)) * kTEdgeRowBytes); };
    double *gin = io.edge_in + 2 * lane - (ptrdiff_t)lvl * kRecDoubles;                       // row consumed by the call for step -1
    double *gout = io.edge_out + 2 * lane - (ptrdiff_t)(2 + kSkew + lvl) * kRecDoubles;      // row sent by that call
    double va = 0.0, vb = 0.0;
    int err = 0;

    if (io.strip > 0) {
        // start distance: not before the strip on the left is kStartRow rows into this time step
        const int rw = kTStartRow < ny ? kTStartRow : ny - 1;
        const double *gw = io.edge_in + (size_t)rw * kRecDoubles + 2 * ((NPR - 1) & ~1);
        SpinClock spins;
        const unsigned long long tw0 = global_timer_ns();
        for (;;) {
            double a = 0.0, b = 0.0;
            if (lane == 0) { if (in_remote) ld_sys2(gw, a, b); else ld_volatile2(gw, a, b); }
            if (__all_sync(0xffffffffu, !(is_sentinel(a) || is_sentinel(b)))) break;
            if (spins.expired()) { err = BFD_ETIMEOUT; break; }
            start_pause();                                   // a waiting warp must not eat the issue slots of the running ones
        }
        if (lane == 0) *io.wait_ns = global_timer_ns() - tw0;
        if (take) {
#pragma unroll
            for (int r = 0; r < kTAhead; ++r)
                if (r < ny) cp_async16(ering_row(r), io.edge_in + 2 * lane + (size_t)r * kRecDoubles);
        }
    }
    // the call for step s: consumes row s + 1 - lvl (into the halo slot read in step s + 1), sends row s - 32 - lvl
    auto halo = [&](const int s) {
        const int rc = s + 1 - lvl, ro = s - 1 - kSkew - lvl;
        const int wbuf = s & 1;
        double na = 0.0, nb = 0.0;
        if (take && (unsigned)(rc + 1) < (unsigned)ny) lds2(ering_row(rc + 1), na, nb);
        if (take && (unsigned)rc < (unsigned)ny) {
            if (is_sentinel_hi(va) || is_sentinel_hi(vb)) {          // had not arrived when the row was copied
                SpinClock spins;
                for (;;) {
                    if (in_remote) ld_sys2(gin, va, vb); else ld_volatile2(gin, va, vb);
                    if (!(is_sentinel(va) || is_sentinel(vb))) break;
                    if (spins.expired()) { err = BFD_ETIMEOUT; va = 0.0; vb = 0.0; break; }
                    if (spins > 2) spin_pause();
                }
            }
            sts2(px + (saddr)(wbuf * kXRow), va, vb);
            st_weak2(gin, sentinel_value(), sentinel_value());       // re-arm for the next launch
            if (rc + kTAhead < ny) cp_async16(ering_row(rc + kTAhead), gin + (size_t)kTAhead * kRecDoubles);
        }
        if (give && (unsigned)ro < (unsigned)ny) {
            double a, b;
            lds2(px + (saddr)((wbuf ^ 1) * kXRow + 32 * 16), a, b);
            if (out_remote) st_sys2(gout, a, b); else st_volatile2(gout, a, b);
        }
        va = na; vb = nb;
        gin += kRecDoubles; gout += kRecDoubles;
    };

    // ---- level-0 input ring ------------------------------------------------------------------
    const saddr ring_lane = io_ring + (saddr)(lane * 8);
    const saddr ring_cp = io_ring + (saddr)(lane * 16);
    const saddr wp_lane = io_xchg + (saddr)(kTWpOffset + lane * 8);
    const saddr wp_cp = io_xchg + (saddr)(kTWpOffset + lane * 16);
    const int in_base = io.in_base, out_base = io.out_base;
    int prod_seen = 0;                                       // cached row counter of the input's producer
    const double *gsrc = io.xin + lane, *gsrcp = io.xwp + lane;   // lane l copies doubles 2l, 2l+1 of a slot-row (u | v: 64 doubles)
    int pf_row = 0;
    auto pf_issue = [&]() {                                  // HBM / L2 -> ring, one slot-row
        if (pf_row <= last_row) {
            if (gated) {
                const int need = in_base + pf_row + 1;
                if (need > prod_seen) {
                    SpinClock spins;
                    while ((prod_seen = ld_shared_volatile(io.prod_in)) < need) {
                        BFD_TRACE_SPIN(spins, "wait prod: strip %d in_base %d row %d need %d seen %d\n", io.strip, in_base, pf_row, need, prod_seen);
                        if (spins.expired()) { err = BFD_ETIMEOUT; break; }
                    }
                    __threadfence_block();
                }
            }
            const int slot = (pf_row & (kTRing - 1)) * 512;
            cp_async16(ring_cp + (saddr)slot, gsrc);
            if (CONT) cp_async16(wp_cp + (saddr)slot, gsrcp);
            gsrc += 2 * W; if (CONT) gsrcp += 2 * W;
            ++pf_row;
        }
    };
#pragma unroll
    for (int i = 0; i < kTPf - 1; ++i) { pf_issue(); cp_async_commit(); }
    cp_async_wait_pending<kTPf - 2>();
    {
        const int rc0 = -lvl;
        if (take && (unsigned)rc0 < (unsigned)ny) lds2(ering_row(rc0), va, vb);
    }
    halo(-1);
    double lb_next = 0.0, cy_next = cyU, cyd_next = cydU;
    if (io.inlet) lb_next = __ldg(io.lbc);
    if (!UNI) {
        cy_next = __ldg(io.cy); cyd_next = __ldg(io.cyd);
    }
    __syncwarp();

    auto body = [&](const int s, auto ep) {
        constexpr bool EP = decltype(ep)::value;
        const int rbuf = (s + 1) & 1, wbuf = s & 1;
        const int srow = s < last_row ? s : last_row;          // slot-row of the level-0 input
        const int so = s - (K - 1);                            // slot-row of x_K written this step
        pf_issue();                                            // row s + kTPf - 1
        cp_async_commit();
        halo(s);
        double up, vp, lb = lb_next;
        {
            const saddr d = ring_lane + (saddr)((srow & (kTRing - 1)) * 512);
            xu[0] = lds(d); xv[0] = lds(d + 256);
            if (CONT) {
                const saddr dp = wp_lane + (saddr)((srow & (kTRing - 1)) * 512);
                up = lds(dp); vp = lds(dp + 256);
            } else { up = xu[0]; vp = xv[0]; }
            if (io.strip == 0) { if (io.inlet) lb_next = __ldg(io.lbc + (s + 1 < ny ? s + 1 : ny - 1)); }   // warp-uniform outer test
            if (!UNI) {
                Cy[0] = cy_next; Cyd[0] = cyd_next;
                const int r = s + 1 - lane;
                const int rcy = r < 0 ? 0 : (r >= ny ? ny - 1 : r);
                cy_next = __ldg(io.cy + rcy); cyd_next = __ldg(io.cyd + rcy);
            }
        }

        // The K+1 levels of a step are independent of one another.  With one loop over the levels the
        // compiler emitted them one after the other (serial DFMA chains, fixed-latency waits 2.2 per
        // issue, ncu r02_tile_v2); so the step is written in two phases: the residuals of ALL levels
        // first (2 (K+1) independent FMA chains side by side), then the K block solves.
        double LF[K + 1], LG[K + 1];
#pragma unroll
        for (int k = K; k >= 0; --k) lds2(xc_left + xc_piece(2 * k, rbuf), LF[k], LG[k]);

        {   // level 0 only: previous-snapshot parts and the level-independent terms Au, Av of row s - lane
            const int rk = s - lane;
            const bool top = EP ? (rk <= 0) : false;
            const double cy = UNI ? cyU : Cy[0], cyd = UNI ? cydU : Cyd[0];
            const double hup = 0.5 * up, hvp = 0.5 * vp;
            const double Fp = hup * up, Gp = hup * vp, Hp = hvp * vp;
            double DGp = CONT ? qGp : prev[0].Gh, DHp = CONT ? qHp : prev[0].Hh;
            if (EP) { DGp = top ? 0.0 : DGp; DHp = top ? 0.0 : DHp; }
            double lFp = LF[0], lGp = LG[0];
            if (CONT) lds2(xc_left + xc_piece(2 * K + 1, rbuf), lFp, lGp);
            double au = fma(cx, Fp, -up);
            au = fma(cy, Gp, au); au = fma(-cxl, lFp, au); au = fma(-cyd, DGp, au);
            au = (au - src) - lb;
            double av = fma(cy, Hp, -vp);
            av = fma(cx, Gp, av); av = fma(-cyd, DHp, av); av = fma(-cxl, lGp, av);
            Au[0] = au; Av[0] = av;
            if (CONT) {
                qGp = Gp; qHp = Hp;
                sts2(xc_own + xc_piece(2 * K + 1, wbuf), Fp, Gp);
            }
        }

        double ru[K + 1], rv[K + 1];
#pragma unroll
        for (int k = K; k >= 0; --k) {                         // phase 1: residual rows of every level
            const int rk = s - lane - k;
            const bool rowok = EP ? (unsigned)rk < (unsigned)ny : true;
            const bool valid = rowok && colok;
            const bool top = EP ? (rk <= 0) : false;
            const double cy = UNI ? cyU : Cy[k], cyd = UNI ? cydU : Cyd[k];
            const double u = xu[k], v = xv[k];
            const double hu = 0.5 * u, hv = 0.5 * v;
            const double Fh = hu * u, Gh = hu * v, Hh = hv * v;
            double DG = prev[k].Gh, DH = prev[k].Hh;
            if (EP) { DG = top ? 0.0 : DG; DH = top ? 0.0 : DH; }
            double a = Au[k] + u;
            a = fma(cx, Fh, a); a = fma(cy, Gh, a); a = fma(-cxl, LF[k], a); a = fma(-cyd, DG, a);
            double b = Av[k] + v;
            b = fma(cy, Hh, b); b = fma(cx, Gh, b); b = fma(-cyd, DH, b); b = fma(-cxl, LG[k], b);
            ru[k] = a; rv[k] = b;
            const double r2 = fma(b, b, a * a);
            nrm[k] += EP ? (valid ? r2 : 0.0) : r2;
            sts2(xc_own + xc_piece(2 * k, wbuf), Fh, Gh);
            prev[k].Gh = Gh; prev[k].Hh = Hh;
        }

        double Lpu[K], Lhpm[K];
#pragma unroll
        for (int k = K - 1; k >= 0; --k) lds2(xc_left + xc_piece(2 * k + 1, rbuf), Lpu[k], Lhpm[k]);
#pragma unroll
        for (int k = K - 1; k >= 0; --k) {                     // phase 2: the 2x2 block solves
            const int rk = s - lane - k;
            const bool rowok = EP ? (unsigned)rk < (unsigned)ny : true;
            const bool valid = rowok && colok;
            const bool top = EP ? (rk <= 0) : false;
            const double cy = UNI ? cyU : Cy[k], cyd = UNI ? cydU : Cyd[k];
            const double u = xu[k], v = xv[k];
            const double hu = 0.5 * u, hv = 0.5 * v;
            double Dhpm = prev[k].pm, Dpvv = prev[k].pvv;        // prev[k].pm holds (v du + u dv) / 2
            if (EP) { Dhpm = top ? 0.0 : Dhpm; Dpvv = top ? 0.0 : Dpvv; }
            const double bu_ = fma(cyd, Dhpm, fma(cxl, Lpu[k], ru[k]));
            const double bv_ = fma(cyd, Dpvv, fma(cxl, Lhpm[k], rv[k]));
            const double d00 = fma(cx, u, fma(cy, hv, 1.0));
            const double d11 = fma(cy, v, fma(cx, hu, 1.0));
            const double d01 = cy * hu, d10 = cx * hv;
            const double inv = rcp_fast(fma(d00, d11, -(d01 * d10)));
            const double du = fma(d11, bu_, -(d01 * bv_)) * inv;
            const double dv = fma(d00, bv_, -(d10 * bu_)) * inv;
            const double pu = u * du, hpm = fma(hv, du, hu * dv);
            prev[k].pm = hpm; prev[k].pvv = v * dv;
            sts2(xc_own + xc_piece(2 * k + 1, wbuf), pu, hpm);
            const double nu = u - du, nv = v - dv;
            if (k == K - 1) {
                if (!EP || (so >= 0 && so <= last_row)) {
                    const double ou = EP ? (valid ? nu : 0.0) : nu, ov = EP ? (valid ? nv : 0.0) : nv;
                    double *o = io.xout + (size_t)so * (2 * W);
                    o[0] = ou; o[W] = ov;
                }
            }
            xu[k + 1] = nu; xv[k + 1] = nv;
            Au[k + 1] = Au[k]; Av[k + 1] = Av[k];
            if (!UNI) { Cy[k + 1] = Cy[k]; Cyd[k + 1] = Cyd[k]; }
        }
        cp_async_wait_pending<kTPf - 2>();
        __syncwarp();
        // rows 0 .. so of x_{t+1} are in the pool: the warp that advances the next time step may fetch them
        if (publish && ((so & (kTPublish - 1)) == kTPublish - 1 || so == last_row) && so >= 0 && so <= last_row) {
            if (lane == 0) { __threadfence_block(); *io.prod_out = out_base + so + 1; }
        }
    };

    const int head_end = nsteps < kSkew + K + 1 ? nsteps : kSkew + K + 1;
    const int main_end = ny > head_end ? ny : head_end;
    int s = 0;
#pragma unroll 1
    for (int phase = 0; phase < 2; ++phase) {
        const int ep_end = phase == 0 ? head_end : nsteps;
#pragma unroll 1
        for (; s < ep_end; ++s) body(s, BoolTag<true>{});
        if (phase == 0) {
#pragma unroll 1
            for (; s < main_end; ++s) body(s, BoolTag<false>{});
        }
    }
    cp_async_commit();
    halo(nsteps);                              // what lane 31 published in the last step
    cp_async_wait_pending<0>();

#pragma unroll
    for (int k = 0; k <= K; ++k) {
        double x = colok ? nrm[k] : 0.0;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) x += __shfl_xor_sync(0xffffffffu, x, off);
        if (lane == 0) io.partial[k] = x;
    }
    if (__any_sync(0xffffffffu, err != 0) && lane == 0) atomicExch(io.error, BFD_ETIMEOUT);
}

template <bool UNI, bool SLAB>
__global__ void __launch_bounds__(32 * kTileDepth, 4) fom_tile_kernel(const FomArgs a) {
    constexpr int D = kTileDepth;
    BFD_SHARED_ARRAY(double, s_mem, [D * kTWarpBytes / 8]);
    BFD_SHARED_ARRAY(int, s_ctl, [8]);
    BFD_SHARED_ARRAY(int, s_prog, [D]);            // rows of x written by warp d in this segment (stream position)
    BFD_SHARED_ARRAY(int, s_end, [2 * D]);         // first attempt (relative) each warp did not run, refused flag
    BFD_SHARED_ARRAY(double, s_dbl, [2]);
    const int lane = lane_id(), warp = threadIdx.x >> 5;
    const int nstrips = a.ssl.nstrips;
    for (int i = threadIdx.x; i < D * kTWarpBytes / 8; i += blockDim.x) s_mem[i] = 0.0;
    if (threadIdx.x == 0) s_ctl[0] = (int)atomicInc(a.ticket, gridDim.x - 1);
    __syncthreads();
    const int item = s_ctl[0];
    if (item >= a.batch * nstrips) return;
    const int m = item / nstrips, j = item - m * nstrips;
    const int jg = j + a.strip0, nsg = a.ns_global;
    FomCtl *ctl = a.ctl + m;
    if (SLAB) {
        if (threadIdx.x == 0 && a.launch_index > 0) {
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
    const int NR = ny + kSkew;                     // slot-rows of a time step = stream positions per chunk

    TileIo io;
    const saddr base = smem_addr(s_mem);
    const int prevw = warp == 0 ? D - 1 : warp - 1;
    io.xchg = base + (saddr)(warp * kTWarpBytes);
    io.ring = io.xchg + (saddr)kTXchgBytes;
    io.edge_ring = io.ring + (saddr)kTRingBytes;
    io.prod_in = (volatile int *)&s_prog[prevw];
    io.prod_out = (volatile int *)&s_prog[warp];
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
        if (threadIdx.x == 0) {
            const volatile FomCtl *vc = ctl;
            s_ctl[1] = vc->t; s_ctl[2] = vc->kbase; s_ctl[3] = vc->k_first; s_ctl[4] = vc->k_spec;
            s_ctl[6] = (int)vc->epoch;
            s_dbl[0] = vc->init_norm;
            for (int d = 0; d < D; ++d) s_prog[d] = 0;
        }
        __syncthreads();
        const int t0 = s_ctl[1];
        if (t0 >= a.num_steps || attempt >= a.chunk) break;
        const int kbase = s_ctl[2], k_first = s_ctl[3], k_spec = s_ctl[4];
        const unsigned epoch0 = (unsigned)s_ctl[6];
        const double init_prev = s_dbl[0];
        const int i0 = attempt;
        int my_end = 0;
        bool refused = false;
        for (int ii = warp;; ii += D) {
            const int att = i0 + ii, t = t0 + ii;
            if (att >= a.chunk || t >= a.num_steps) { my_end = ii; break; }
            int ok = 1;
            if (lane == 0) {
                const unsigned old = a_max<SLAB>(&ctl->gate, (unsigned)att);
                ok = ((old & kGateStop) && (unsigned)att > (old & (kGateStop - 1))) ? 0 : 1;
            }
            ok = __shfl_sync(0xffffffffu, ok, 0);
            if (!ok) { refused = true; my_end = ii; break; }

            const int kb = ii ? 0 : kbase;
            int K = ii ? k_spec : k_first;
            if (K > a.max_its - kb) K = a.max_its - kb;
            if (K > kTileKMax) K = kTileKMax;
            io.xwp = pool_m + (size_t)t * sd;
            io.xout = pool_m + (size_t)(t + 1) * sd;
            io.xin = kb ? io.xout : io.xwp;
            const size_t eoff = (size_t)att * ny * kRecDoubles;
            io.edge_in = edge_m + (size_t)(j > 0 ? j - 1 : 0) * a.edge_stride_strip + eoff;
            io.edge_out = edge_m + (size_t)(j < nstrips - 1 ? j : 0) * a.edge_stride_strip + eoff;
            if (SLAB) {
                if (io.in_remote) io.edge_in = a.halo + (size_t)m * a.edge_stride_strip + eoff;
                if (io.out_remote) io.edge_out = a.next_halo + (size_t)m * a.edge_stride_strip + eoff;
            }
            io.partial = a.partial + ((size_t)(m * a.chunk + att) * nsg + jg) * kNormSlots;
            io.gated = ii > 0;                                   // attempt ii - 1 (warp prevw) writes this step's input
            io.publish = att + 1 < a.chunk && t + 1 < a.num_steps;
            io.in_base = ii > 0 ? ((ii - 1) / D) * NR : 0;
            io.out_base = (ii / D) * NR;
            unsigned long long *tr = a.trace + ((size_t)(m * a.chunk + att) * nstrips + j) * 3;
            if (lane == 0) { tr[0] = global_timer_ns(); tr[2] = 0; }
            io.wait_ns = tr + 2;
            if (kb) {
                tile_step<1, true, UNI>(io);
            } else {
                switch (K) {
                case 1: tile_step<1, false, UNI>(io); break;
                case 2: tile_step<2, false, UNI>(io); break;
                case 3: tile_step<3, false, UNI>(io); break;
                default: tile_step<4, false, UNI>(io); break;
                }
            }
            __syncwarp();
            if (lane == 0) {
                tr[1] = global_timer_ns();
                fence<SLAB>();
                if (a_inc<SLAB>(&a.arrive[(size_t)m * a.chunk + att], (unsigned)nsg - 1) == (unsigned)nsg - 1) {
                    fence<SLAB>();
                    if (validate_step(a, m, att, kb, kb ? 1 : K, init_prev)) a_or<SLAB>(&ctl->gate, kGateStop);
                }
            }
            __syncwarp();
        }
        if (lane == 0) { s_end[warp] = my_end; s_end[D + warp] = refused ? 1 : 0; }
        __syncthreads();
        int n_run = s_end[0], was_refused = s_end[D];
#pragma unroll
        for (int d = 1; d < D; ++d)
            if (s_end[d] < n_run) { n_run = s_end[d]; was_refused = s_end[D + d]; }
        attempt = i0 + n_run;
        if (threadIdx.x == 0) {
            fence<SLAB>();
            if (a_inc<SLAB>(&ctl->arrive, (unsigned)nsg - 1) == (unsigned)nsg - 1) {
                fence<SLAB>();
                commit_segment(a, m, t0, i0, n_run, kbase, k_first, k_spec);
                *(volatile unsigned *)&ctl->gate = was_refused ? (unsigned)(attempt > 0 ? attempt - 1 : 0) : 0u;
                fence<SLAB>();
                a_add<SLAB>(&ctl->epoch, 1u);
            } else if (was_refused) {
                SpinClock spins;
                while (*(volatile unsigned *)&ctl->epoch == epoch0) {
                    spin_pause();
                    BFD_TRACE_SPIN(spins, "wait epoch: strip %d attempt %d\n", jg, attempt);
                    if (spins.expired()) { atomicExch(&a.glob->error, BFD_ETIMEOUT); break; }
                }
                fence<SLAB>();
            }
        }
        if (!was_refused) break;
    }
    if (SLAB) {
        __syncthreads();
        if (threadIdx.x == 0) { fence<SLAB>(); a_add<SLAB>(&ctl->exits, 1u); }
    }
}

}  // namespace bfd
