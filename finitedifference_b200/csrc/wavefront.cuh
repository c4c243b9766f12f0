// Device building blocks of the wavefront kernels: the per-cell Newton update and the
// strip-to-strip hand-off.
#pragma once
#include "common.cuh"

namespace bfd {

__device__ __forceinline__ double shfl_up1(double x) { return __shfl_up_sync(0xffffffffu, x, 1); }

// 16-byte hand-off accesses: one transaction each on the GPU, so a pair is seen whole.
#ifndef BFD_EMULATE
__device__ __forceinline__ void ld_volatile2(const double *p, double &a, double &b) {
    asm volatile("ld.volatile.global.v2.f64 {%0, %1}, [%2];" : "=d"(a), "=d"(b) : "l"(p) : "memory");
}
__device__ __forceinline__ void st_volatile2(double *p, double a, double b) {
    asm volatile("st.volatile.global.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(a), "d"(b) : "memory");
}
#else   // emulation: second half first / first half last, so the checked half implies the other
inline void ld_volatile2(const double *p, double &a, double &b) {
    std::this_thread::yield();
    unsigned long long x = __atomic_load_n((const unsigned long long *)p, __ATOMIC_ACQUIRE);
    unsigned long long y = __atomic_load_n((const unsigned long long *)p + 1, __ATOMIC_ACQUIRE);
    memcpy(&a, &x, 8); memcpy(&b, &y, 8);
}
inline void st_volatile2(double *p, double a, double b) {
    unsigned long long x, y; memcpy(&x, &a, 8); memcpy(&y, &b, 8);
    __atomic_store_n((unsigned long long *)p + 1, y, __ATOMIC_RELEASE);
    __atomic_store_n((unsigned long long *)p, x, __ATOMIC_RELEASE);
}
#endif
__device__ __forceinline__ bool is_sentinel(double x) {
    return (unsigned long long)__double_as_longlong(x) == kSentinelBits;
}
__device__ __forceinline__ double sentinel_value() { return __longlong_as_double((long long)kSentinelBits); }

// What a cell hands to its right (L role) and lower (D role) neighbours.
struct Nbr {
    double u, v;    // state x_k
    double F, G, H; // trapezoid flux sums (u^2+up^2)/2, (uv+up vp)/2, (v^2+vp^2)/2
    double du, dv;  // Newton correction
};
__device__ __forceinline__ Nbr nbr_zero() { return Nbr{0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0}; }

// Fused residual (hypernet2D.py:2619-2643) + one block row of the forward substitution with the
// 2x2 blocks of inviscid_burgers_exact_jac2D (:2719-2732):
//     D_p d_p = r_p - A_L d_L - A_D d_D
//   D_p = [[1 + cx u + cy v/2, cy u/2], [cx v/2, 1 + cy v + cx u/2]]
//   A_L = -cxl [[u_L, 0], [v_L/2, u_L/2]]      A_D = -cyd [[v_D/2, u_D/2], [0, v_D]]
// `srcl` = source term (+ inlet ghost flux on column 0).  Returns the cell's Nbr record and the
// squared residual; the caller masks cells outside the grid.
__device__ __forceinline__ Nbr cell_update(double u, double v, double up, double vp, double cx, double cxl,
                                           double cy, double cyd, double src, double lb, const Nbr &L,
                                           const Nbr &D, double &r2) {
    Nbr o;
    o.u = u; o.v = v;
    o.F = 0.5 * (u * u) + 0.5 * (up * up);
    o.H = 0.5 * (v * v) + 0.5 * (vp * vp);
    o.G = 0.5 * u * v + 0.5 * up * vp;
    const double ru = ((u - up) + (cx * o.F - cxl * L.F) + (cy * o.G - cyd * D.G) - src) - lb;
    const double rv = (v - vp) + (cy * o.H - cyd * D.H) + (cx * o.G - cxl * L.G);
    r2 = ru * ru + rv * rv;
    const double bu = ru + cxl * (L.u * L.du) + 0.5 * cyd * (D.v * D.du + D.u * D.dv);
    const double bv = rv + 0.5 * cxl * (L.v * L.du + L.u * L.dv) + cyd * (D.v * D.dv);
    const double d00 = 1.0 + cx * u + 0.5 * cy * v;
    const double d01 = 0.5 * cy * u;
    const double d10 = 0.5 * cx * v;
    const double d11 = 1.0 + cy * v + 0.5 * cx * u;
    const double inv = 1.0 / (d00 * d11 - d01 * d10);
    o.du = (d11 * bu - d01 * bv) * inv;
    o.dv = (d00 * bv - d10 * bu) * inv;
    return o;
}

// ------------------------------------------------------------------------------------------
// Strip-to-strip hand-off.  The last lane of a strip produces, per row, the record
// {u, v, F, G, du, dv} its right neighbour strip needs (48 bytes = three 16-byte pieces).
// Records travel in blocks of kEdgeBlock rows: the producer collects them in shared memory and
// the warp flushes a block with coalesced 16-byte volatile stores; the consumer warp fetches a
// block the same way one block ahead, parks it in shared memory and its lane 0 reads one
// record per step.  No flags, no fences: a piece is valid once its first double is not the
// sentinel (the buffer starts armed, and the consumer re-arms what it has taken, ready for the
// next sweep).
// ------------------------------------------------------------------------------------------
constexpr int kEdgeBlock = 8;
constexpr int kEdgePieces = kEdgeBlock * 3;          // 16-byte pieces per block (<= 32)
constexpr int kEdgeSmemDoubles = kEdgeBlock * kEdgeDoubles;
constexpr unsigned kSpinLimit = 1u << 22;

struct EdgeReader {
    double *g;       // records of the strip on the left, [ny][kEdgeDoubles]
    double *sm;      // kEdgeSmemDoubles doubles of shared memory owned by this warp
    double p0, p1;   // piece fetched ahead
    int ny, lane;
    bool active, dead;

    __device__ __forceinline__ void init(double *g_, double *sm_, int ny_, int lane_, bool active_) {
        g = g_; sm = sm_; ny = ny_; lane = lane_; active = active_; dead = false;
        p0 = p1 = 0.0;
        if (active) fetch(0);
    }
    __device__ __forceinline__ bool mine(int row0) const { return lane < kEdgePieces && row0 + lane / 3 < ny; }
    __device__ __forceinline__ void fetch(int row0) {
        if (mine(row0)) ld_volatile2(g + (size_t)row0 * kEdgeDoubles + lane * 2, p0, p1);
    }
    // Call by the whole warp at every step; does work when `s` opens a block.  Returns an
    // error code (BFD_ETIMEOUT) once if the producer never showed up.
    __device__ __forceinline__ int step(int s) {
        if (!active || s >= ny || (s % kEdgeBlock) != 0) return 0;
        int err = 0;
        if (!dead) {
            unsigned spins = 0;
            for (;;) {
                const bool ok = !mine(s) || !is_sentinel(p0);
                if (__all_sync(0xffffffffu, ok)) break;
                if (!ok) ld_volatile2(g + (size_t)s * kEdgeDoubles + lane * 2, p0, p1);
                if (++spins > kSpinLimit) { dead = true; err = BFD_ETIMEOUT; break; }
            }
        }
        __syncwarp();                      // lane 0 is done with the previous block
        if (mine(s)) {
            sm[lane * 2] = dead ? 0.0 : p0;
            sm[lane * 2 + 1] = dead ? 0.0 : p1;
            if (!dead) st_volatile2(g + (size_t)s * kEdgeDoubles + lane * 2, sentinel_value(), sentinel_value());
        }
        __syncwarp();
        if (!dead && s + kEdgeBlock < ny) fetch(s + kEdgeBlock);
        return err;
    }
    // lane 0: the record of row s (u, v, F, G, du, dv)
    __device__ __forceinline__ void get(int s, Nbr &L) const {
        const double *r = sm + (s % kEdgeBlock) * kEdgeDoubles;
        L.u = r[0]; L.v = r[1]; L.F = r[2]; L.G = r[3]; L.du = r[4]; L.dv = r[5];
        L.H = 0.0;
    }
};

struct EdgeWriter {
    double *g;   // records of this strip
    double *sm;
    int ny, lane;
    bool active;

    __device__ __forceinline__ void init(double *g_, double *sm_, int ny_, int lane_, bool active_) {
        g = g_; sm = sm_; ny = ny_; lane = lane_; active = active_;
    }
    // Call by the whole warp at every step with the row `r31` lane 31 has just finished and its
    // record (only lane 31's `c` is used).
    __device__ __forceinline__ void step(int r31, const Nbr &c) {
        if (!active || r31 < 0 || r31 >= ny) return;
        if (lane == 31) {
            double *r = sm + (r31 % kEdgeBlock) * kEdgeDoubles;
            r[0] = c.u; r[1] = c.v; r[2] = c.F; r[3] = c.G; r[4] = c.du; r[5] = c.dv;
        }
        if ((r31 % kEdgeBlock) != kEdgeBlock - 1 && r31 != ny - 1) return;
        const int row0 = r31 - (r31 % kEdgeBlock);
        __syncwarp();
        if (lane < kEdgePieces && row0 + lane / 3 < ny)
            st_volatile2(g + (size_t)row0 * kEdgeDoubles + lane * 2, sm[lane * 2], sm[lane * 2 + 1]);
        __syncwarp();
    }
};

}  // namespace bfd
