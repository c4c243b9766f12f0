// Device building blocks of the fused wavefront kernel: small PTX wrappers and the
// strip-to-strip hand-off of per-row, per-Newton-level records.
#pragma once
#include "common.cuh"

namespace bfd {

__device__ __forceinline__ double shfl_up1(double x) { return __shfl_up_sync(0xffffffffu, x, 1); }

#ifndef BFD_EMULATE
// 16-byte hand-off accesses (one transaction each), relaxed at GPU scope: coherent in L2,
// without the system-scope ordering of ld/st.volatile (which serialises back-to-back stores).
__device__ __forceinline__ void ld_volatile2(const double *p, double &a, double &b) {
    asm volatile("ld.relaxed.gpu.global.v2.f64 {%0, %1}, [%2];" : "=d"(a), "=d"(b) : "l"(p) : "memory");
}
__device__ __forceinline__ void st_volatile2(double *p, double a, double b) {
    asm volatile("st.relaxed.gpu.global.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(a), "d"(b) : "memory");
}
// plain 16-byte store: re-arming a record only has to be visible to the next launch
__device__ __forceinline__ void st_weak2(double *p, double a, double b) {
    asm volatile("st.global.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(a), "d"(b) : "memory");
}
// system scope: records that cross NVLink into (or are polled from under) a peer GPU's writes
__device__ __forceinline__ void ld_sys2(const double *p, double &a, double &b) {
    asm volatile("ld.relaxed.sys.global.v2.f64 {%0, %1}, [%2];" : "=d"(a), "=d"(b) : "l"(p) : "memory");
}
__device__ __forceinline__ void st_sys2(double *p, double a, double b) {
    asm volatile("st.relaxed.sys.global.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(a), "d"(b) : "memory");
}
// Shared memory is addressed by 32-bit shared-window byte addresses computed once per kernel,
// so the hot loop never re-derives a generic pointer (S2R + LEA per access otherwise).
typedef unsigned saddr;
__device__ __forceinline__ saddr smem_addr(const void *p) { return (saddr)__cvta_generic_to_shared(p); }
__device__ __forceinline__ double lds(saddr a) {
    double v; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a) : "memory"); return v;
}
__device__ __forceinline__ void lds2(saddr a, double &x, double &y) {
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x), "=d"(y) : "r"(a) : "memory");
}
__device__ __forceinline__ void sts(saddr a, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v) : "memory"); }
__device__ __forceinline__ void sts2(saddr a, double x, double y) {
    asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(a), "d"(x), "d"(y) : "memory");
}
__device__ __forceinline__ int lane_id() { int l; asm volatile("mov.u32 %0, %%laneid;" : "=r"(l)); return l; }
// 16 bytes global -> shared without a register round trip (LDGSTS, L2 only)
__device__ __forceinline__ void cp_async16(saddr dst, const double *gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(gsrc) : "memory");
}
// 8 bytes global -> shared (each lane copies its own element)
__device__ __forceinline__ void cp_async8(saddr dst, const double *gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(gsrc) : "memory");
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
typedef uintptr_t saddr;
inline saddr smem_addr(const void *p) { return (saddr)p; }
inline double lds(saddr a) { return *(const double *)a; }
inline void lds2(saddr a, double &x, double &y) { x = ((const double *)a)[0]; y = ((const double *)a)[1]; }
inline void sts(saddr a, double v) { *(double *)a = v; }
inline void sts2(saddr a, double x, double y) { ((double *)a)[0] = x; ((double *)a)[1] = y; }
inline int lane_id() { return emu::tls.lane; }
inline void cp_async16(saddr dst, const double *src) { ld_volatile2(src, ((double *)dst)[0], ((double *)dst)[1]); }
inline void cp_async8(saddr dst, const double *src) { *(double *)dst = *src; }
inline void st_weak2(double *p, double a, double b) { st_volatile2(p, a, b); }
inline void ld_sys2(const double *p, double &a, double &b) { ld_volatile2(p, a, b); }
inline void st_sys2(double *p, double a, double b) { st_volatile2(p, a, b); }
inline void cp_async_commit() {}
template <int N> inline void cp_async_wait_pending() {}
inline double rcp_fast(double x) { return 1.0 / x; }
inline void spin_pause() { std::this_thread::yield(); }
inline unsigned long long global_timer_ns() { return 0; }
#endif

// keep_reg: stops the compiler from rematerialising a loop-invariant value (shared-window addresses cost an
// S2R + LEA, constants a MOV each) in every wavefront step: the value stays in the register it is in.
#ifndef BFD_EMULATE
__device__ __forceinline__ unsigned keep_reg(unsigned v) { asm volatile("" : "+r"(v)); return v; }
__device__ __forceinline__ int keep_reg(int v) { asm volatile("" : "+r"(v)); return v; }
__device__ __forceinline__ double keep_reg(double v) { asm volatile("" : "+d"(v)); return v; }
#else
inline saddr keep_reg(saddr v) { return v; }
inline int keep_reg(int v) { return v; }
inline double keep_reg(double v) { return v; }
#endif

__device__ __forceinline__ bool is_sentinel(double x) {
    return (unsigned long long)__double_as_longlong(x) == kSentinelBits;
}
// cheaper test for values that came through the ring: no arithmetic result has these high 32 bits
__device__ __forceinline__ bool is_sentinel_hi(double x) {
    return (unsigned)((unsigned long long)__double_as_longlong(x) >> 32) == (unsigned)(kSentinelBits >> 32);
}
__device__ __forceinline__ double sentinel_value() { return __longlong_as_double((long long)kSentinelBits); }

// ------------------------------------------------------------------------------------------
// Strip-to-strip hand-off.  For every grid row the last lane of a strip produces what the
// first column of the strip on its right needs from its left neighbour, for every Newton level
// of the fused pass: NPR 16-byte pieces per row
//     piece 2k   = (Fh_k, Gh_k)      current-iterate flux parts of level k  (k = 0..K)
//     piece 2k+1 = (pu_k, pm_k)      u*du and (v*du + u*dv)/2 of level k    (k < K)
//     piece 2K+1 = (Fp, Gp)          previous-time-level flux parts (continuation passes only)
// The producer's data-movement warp stores each piece the step after it was computed (one
// 16-byte relaxed store per lane); the consumer's copies a block of kEdgeBlock rows
// into its shared-memory ring one block ahead (cp.async, in the caller's per-step commit
// groups: the caller waits for the group before open()), checks it when the block opens and
// spins on whatever had not arrived.  No flags, no fences: a piece is valid once neither of its
// doubles is the sentinel (the buffer starts armed and the consumer re-arms what it has taken,
// ready for the next launch).
// ------------------------------------------------------------------------------------------
constexpr int kEdgeBlock = 8;
constexpr int kPiecesMax = 2 * BFD_FOM_KMAX + 2;   // pieces per row in global memory (fixed pitch)
constexpr int kRecDoubles = 2 * kPiecesMax;        // doubles per row record
constexpr int kInRing = 32;                        // rows in the consumer's ring
constexpr unsigned kSpinLimit = 1u << 22;                       // emulator: iterations
constexpr unsigned long long kSpinTimeoutNs = 20000000000ull;   // device: 20 s of %globaltimer

// Deadline of a device-side wait.  A wait gives up (BFD_ETIMEOUT) after kSpinTimeoutNs of wall time, not
// after a number of iterations: how long an iteration takes depends on where the polled word lives (L2,
// a peer GPU over NVLink) and on what else runs on the SM, and in multi-process slab runs the neighbour's
// kernel is launched by another process (launch skew of seconds is possible on a busy host).  The timer
// is read every 1024 iterations only.
struct SpinClock {
    unsigned n = 0;
    unsigned long long t0 = 0;
    __device__ __forceinline__ bool expired() {
#ifdef BFD_EMULATE
        return ++n > kSpinLimit;
#else
        if ((++n & 1023u) != 0) return false;
        const unsigned long long t = global_timer_ns();
        if (t0 == 0) { t0 = t; return false; }
        return t - t0 > kSpinTimeoutNs;
#endif
    }
    __device__ __forceinline__ operator unsigned() const { return n; }
};

constexpr int kRecBytes = kRecDoubles * 8;

constexpr int kAhead = 12;                // rows copied into the consumer's ring ahead of their use
constexpr int kStartRow = kAhead + 6;     // a strip starts a time step when its producer has finished this row
static_assert(kAhead + 2 <= kInRing, "rows in flight plus the row in use must fit the ring");

}  // namespace bfd
