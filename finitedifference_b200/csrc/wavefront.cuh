// Device building blocks of the fused wavefront kernel: small PTX wrappers and the
// strip-to-strip hand-off of per-row, per-Newton-level records.
#pragma once
#include "common.cuh"

namespace bfd {

__device__ __forceinline__ double shfl_up1(double x) { return __shfl_up_sync(0xffffffffu, x, 1); }

#ifndef BFD_EMULATE
// 16-byte hand-off accesses (one transaction each, L2-coherent).
__device__ __forceinline__ void ld_volatile2(const double *p, double &a, double &b) {
    asm volatile("ld.volatile.global.v2.f64 {%0, %1}, [%2];" : "=d"(a), "=d"(b) : "l"(p) : "memory");
}
__device__ __forceinline__ void st_volatile2(double *p, double a, double b) {
    asm volatile("st.volatile.global.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(a), "d"(b) : "memory");
}
// 16 bytes global -> shared without a register round trip (LDGSTS, L2 only)
__device__ __forceinline__ void cp_async16(double *smem_dst, const double *gsrc) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gsrc) : "memory");
}
// 8 bytes global -> shared (each lane copies its own element)
__device__ __forceinline__ void cp_async8(double *smem_dst, const double *gsrc) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
// wait until at most N of this thread's committed groups are still in flight
template <int N>
__device__ __forceinline__ void cp_async_wait_pending() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
// 1/x for x of order one: 20-bit seed + two Newton steps (relative error ~1e-16; the Newton
// correction it scales tolerates far more).  5 FP64-pipe instructions instead of ~10.
__device__ __forceinline__ double rcp_fast(double x) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double e = fma(-x, r, 1.0);
    r = fma(r, e, r);
    e = fma(-x, r, 1.0);
    return fma(r, e, r);
}
__device__ __forceinline__ void spin_pause() { __nanosleep(20); }
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#else
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
inline void cp_async16(double *dst, const double *src) { ld_volatile2(src, dst[0], dst[1]); }
inline void cp_async8(double *dst, const double *src) { *dst = *src; }
inline void cp_async_commit() {}
template <int N> inline void cp_async_wait_pending() {}
inline double rcp_fast(double x) { return 1.0 / x; }
inline void spin_pause() { std::this_thread::yield(); }
inline unsigned long long global_timer_ns() { return 0; }
#endif

__device__ __forceinline__ bool is_sentinel(double x) {
    return (unsigned long long)__double_as_longlong(x) == kSentinelBits;
}
__device__ __forceinline__ double sentinel_value() { return __longlong_as_double((long long)kSentinelBits); }

// ------------------------------------------------------------------------------------------
// Strip-to-strip hand-off.  For every grid row the last lane of a strip produces what the
// first column of the strip on its right needs from its left neighbour, for every Newton level
// of the fused pass: NPR 16-byte pieces per row
//     piece 2k   = (Fh_k, Gh_k)      current-iterate flux parts of level k  (k = 0..K)
//     piece 2k+1 = (pu_k, pm_k)      u*du and v*du + u*dv of level k        (k < K)
//     piece 2K+1 = (Fp, Gp)          previous-time-level flux parts (continuation passes only)
// Rows travel in blocks of kEdgeBlock: the producer collects them in shared memory and the
// warp flushes a block with coalesced 16-byte volatile stores; the consumer warp copies a block
// into its shared-memory ring one block ahead (cp.async, in the caller's per-step commit
// groups: the caller waits for the group before open()), checks it when the block opens and
// spins on whatever had not arrived.  No flags, no fences: a piece is valid once neither of its
// doubles is the sentinel (the buffer starts armed and the consumer re-arms what it has taken,
// ready for the next launch).
// ------------------------------------------------------------------------------------------
constexpr int kEdgeBlock = 8;
constexpr int kPiecesMax = 2 * BFD_FOM_KMAX + 2;   // pieces per row in global memory (fixed pitch)
constexpr int kRecDoubles = 2 * kPiecesMax;        // doubles per row record
constexpr int kInRing = 32;                        // rows in the consumer's ring (4 blocks)
constexpr int kOutRing = 16;                       // rows in the producer's ring (2 blocks)
constexpr unsigned kSpinLimit = 1u << 22;

template <int NPR>
struct EdgeIn {
    static constexpr int kIters = (kEdgeBlock * NPR + 31) / 32;
    double *g;       // records written by the strip on the left, [ny][kRecDoubles]
    double *ring;    // [kInRing][kRecDoubles] shared memory of this warp
    int ny, lane;
    bool active, dead;

    // `opener`: this warp is the one that calls open() (the others only read rows)
    __device__ __forceinline__ void init(double *g_, double *ring_, int ny_, int lane_, bool active_, bool opener) {
        g = g_; ring = ring_; ny = ny_; lane = lane_; active = active_; dead = false;
        if (active && opener) prefetch(0);
    }
    // start copying the block whose first row is row0 into the ring
    __device__ __forceinline__ void prefetch(int row0) {
#pragma unroll
        for (int it = 0; it < kIters; ++it) {
            const int q = lane + 32 * it, row = row0 + q / NPR, piece = q % NPR;
            if (q < kEdgeBlock * NPR && row < ny)
                cp_async16(ring + (row % kInRing) * kRecDoubles + 2 * piece, g + (size_t)row * kRecDoubles + 2 * piece);
        }
    }
    // Whole warp, every step; works when step `s` opens a block.  Returns BFD_ETIMEOUT once if
    // the producer never showed up (the block then reads as zeros).  Every branch is
    // warp-uniform (votes), so the caller's shuffles stay convergent.
    __device__ __forceinline__ int open(int s) {
        if (!active || s >= ny || (s % kEdgeBlock) != 0) return 0;
        int err = 0;
        double p0[kIters], p1[kIters];
        bool mine[kIters];
#pragma unroll
        for (int it = 0; it < kIters; ++it) {
            const int q = lane + 32 * it, row = s + q / NPR, piece = q % NPR;
            mine[it] = q < kEdgeBlock * NPR && row < ny;
            const double *slot = ring + ((mine[it] ? row : s) % kInRing) * kRecDoubles + 2 * (mine[it] ? piece : 0);
            p0[it] = slot[0]; p1[it] = slot[1];
        }
        if (!dead) {
            unsigned spins = 0;
            for (;;) {
                bool ok = true;
#pragma unroll
                for (int it = 0; it < kIters; ++it)
                    ok = ok && (!mine[it] || !(is_sentinel(p0[it]) || is_sentinel(p1[it])));
                if (__all_sync(0xffffffffu, ok)) break;
                if (++spins > kSpinLimit) { dead = true; err = BFD_ETIMEOUT; break; }
                if (spins > 2) spin_pause();
#pragma unroll
                for (int it = 0; it < kIters; ++it) {
                    const int q = lane + 32 * it, row = s + q / NPR, piece = q % NPR;
                    if (mine[it] && (is_sentinel(p0[it]) || is_sentinel(p1[it])))
                        ld_volatile2(g + (size_t)row * kRecDoubles + 2 * piece, p0[it], p1[it]);
                }
            }
        }
#pragma unroll
        for (int it = 0; it < kIters; ++it) {
            const int q = lane + 32 * it, row = s + q / NPR, piece = q % NPR;
            if (mine[it]) {
                double *slot = ring + (row % kInRing) * kRecDoubles + 2 * piece;
                slot[0] = dead ? 0.0 : p0[it];
                slot[1] = dead ? 0.0 : p1[it];
                if (!dead) st_volatile2(g + (size_t)row * kRecDoubles + 2 * piece, sentinel_value(), sentinel_value());   // re-arm
            }
        }
        __syncwarp();
        if (!dead && s + kEdgeBlock < ny) prefetch(s + kEdgeBlock);
        return err;
    }
    __device__ __forceinline__ const double *row(int r) const { return ring + (r % kInRing) * kRecDoubles; }
};

template <int NPR>
struct EdgeOut {
    static constexpr int kIters = (kEdgeBlock * NPR + 31) / 32;
    double *g;      // records of this strip, [ny][kRecDoubles]
    double *ring;   // [kOutRing][kRecDoubles]
    int ny, lane;
    bool active;

    __device__ __forceinline__ void init(double *g_, double *ring_, int ny_, int lane_, bool active_) {
        g = g_; ring = ring_; ny = ny_; lane = lane_; active = active_;
    }
    __device__ __forceinline__ double *row(int r) const { return ring + (r % kOutRing) * kRecDoubles; }
    // Whole warp, every step, with the row whose LAST level lane 31 has just finished.
    __device__ __forceinline__ void close(int r_done) {
        if (!active || r_done < 0 || r_done >= ny) return;
        if ((r_done % kEdgeBlock) != kEdgeBlock - 1 && r_done != ny - 1) return;
        const int row0 = r_done - (r_done % kEdgeBlock);
        __syncwarp();
#pragma unroll
        for (int it = 0; it < kIters; ++it) {
            const int q = lane + 32 * it, row = row0 + q / NPR, piece = q % NPR;
            if (q < kEdgeBlock * NPR && row < ny) {
                const double *slot = ring + (row % kOutRing) * kRecDoubles + 2 * piece;
                st_volatile2(g + (size_t)row * kRecDoubles + 2 * piece, slot[0], slot[1]);
            }
        }
        __syncwarp();
    }
};

}  // namespace bfd
