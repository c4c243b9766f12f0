// Minimal SIMT emulator -- TEST INFRASTRUCTURE ONLY.
//
// Compiles the product's .cu sources with g++ (-DBFD_EMULATE -x c++) so that the kernel
// logic -- skewed-strip indexing, warp shuffles, strip hand-off records, the device-side
// Newton bookkeeping -- can be exercised by the CPU test-suite in the build container, where
// there is no GPU.  Every CUDA thread is a host thread; a warp's lanes meet at a barrier for
// each *_sync collective, a CTA's threads at __syncthreads().  It is never linked into
// finitedifference_b200/_burgers_b200.so and the package cannot load it.
#pragma once
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <functional>
#include <memory>
#include <thread>
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __noinline__
#define __launch_bounds__(...)
#define __restrict__
#ifndef __align__
#define __align__(n) __attribute__((aligned(n)))
#endif

struct dim3 {
    unsigned x, y, z;
    dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct double2 { double x, y; };
static inline double2 make_double2(double a, double b) { return double2{a, b}; }

typedef int cudaError_t;
typedef void *cudaStream_t;
typedef void *cudaEvent_t;
enum { cudaSuccess = 0 };
enum { cudaMemcpyDeviceToHost = 0, cudaMemcpyDeviceToDevice = 1, cudaMemcpyHostToDevice = 2 };
enum { cudaEventDisableTiming = 2 };
static inline const char *cudaGetErrorName(cudaError_t) { return "emu"; }
static inline const char *cudaGetErrorString(cudaError_t) { return "emu"; }
static inline cudaError_t cudaGetLastError() { return cudaSuccess; }
static inline cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, int, cudaStream_t) { memcpy(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, int) { memcpy(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemsetAsync(void *d, int v, size_t n, cudaStream_t) { memset(d, v, n); return cudaSuccess; }
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t *e, int) { *e = nullptr; return cudaSuccess; }
static inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaEventDestroy(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaGetDevice(int *d) { *d = 0; return cudaSuccess; }
// page-locked memory is "mapped" unless a test says otherwise (bfd_fom_run then polls through copies)
inline int &emu_host_unmapped_flag() { static int v = 0; return v; }
extern "C" __attribute__((used, weak)) void emu_set_host_unmapped(int v) { emu_host_unmapped_flag() = v; }
static inline cudaError_t cudaHostGetDevicePointer(void **d, void *h, unsigned) {
    if (emu_host_unmapped_flag()) return (cudaError_t)1;
    *d = h;
    return cudaSuccess;
}
// a "device" array is a host array here, so its address is the symbol's own
template <class T> static inline cudaError_t cudaGetSymbolAddress(void **p, T &sym) { *p = (void *)&sym; return cudaSuccess; }

namespace emu {

struct Barrier {  // sense-reversing, yielding (threads outnumber cores)
    std::atomic<int> count{0}, sense{0};
    int n = 0;
    void wait() {
        const int s = sense.load(std::memory_order_acquire);
        if (count.fetch_add(1, std::memory_order_acq_rel) == n - 1) {
            count.store(0, std::memory_order_relaxed);
            sense.store(s ^ 1, std::memory_order_release);
        } else {
            while (sense.load(std::memory_order_acquire) == s) std::this_thread::yield();
        }
    }
};

struct Warp {
    Barrier bar;
    unsigned long long slot[32];
};

struct Cta {
    Barrier bar;
    std::vector<std::unique_ptr<Warp>> warps;
    std::vector<std::vector<unsigned char>> shared;  // one arena per __shared__ declaration site
    std::atomic<int> shared_lock{0};
    void *get_shared(int id, size_t bytes) {
        while (shared_lock.exchange(1, std::memory_order_acquire)) std::this_thread::yield();
        if ((int)shared.size() <= id) shared.resize(id + 1);
        if (shared[id].empty()) shared[id].assign(bytes, 0);
        void *p = shared[id].data();
        shared_lock.store(0, std::memory_order_release);
        return p;
    }
};

struct ThreadCtx {
    dim3 threadIdx, blockIdx, blockDim, gridDim;
    Cta *cta = nullptr;
    Warp *warp = nullptr;
    int lane = 0;
};
inline thread_local ThreadCtx tls;

// Run a grid.  coresident = true starts every CTA at once (kernels whose CTAs wait on one
// another); otherwise CTAs run one after the other.
template <class F>
void launch(dim3 grid, dim3 block, bool coresident, F body) {
    const unsigned nthreads = block.x * block.y * block.z;
    const unsigned nctas = grid.x * grid.y * grid.z;
    const unsigned nwarps = (nthreads + 31) / 32;
    auto run_ctas = [&](unsigned first, unsigned count) {
        std::vector<std::unique_ptr<Cta>> ctas;
        std::vector<std::thread> threads;
        for (unsigned b = first; b < first + count; ++b) {
            ctas.emplace_back(new Cta);
            Cta *cta = ctas.back().get();
            cta->bar.n = (int)nthreads;
            for (unsigned w = 0; w < nwarps; ++w) {
                cta->warps.emplace_back(new Warp);
                const unsigned lanes = (w == nwarps - 1 && nthreads % 32) ? nthreads % 32 : 32;
                cta->warps.back()->bar.n = (int)lanes;
            }
            for (unsigned t = 0; t < nthreads; ++t) {
                threads.emplace_back([=, &body] {
                    ThreadCtx &c = tls;
                    c.blockDim = block; c.gridDim = grid;
                    c.blockIdx = dim3(b % grid.x, (b / grid.x) % grid.y, b / (grid.x * grid.y));
                    c.threadIdx = dim3(t % block.x, (t / block.x) % block.y, t / (block.x * block.y));
                    c.cta = cta; c.warp = cta->warps[t / 32].get(); c.lane = (int)(t % 32);
                    body();
                });
            }
        }
        for (auto &th : threads) th.join();
    };
    if (coresident) run_ctas(0, nctas);
    else for (unsigned b = 0; b < nctas; ++b) run_ctas(b, 1);
}

template <class T>
T shfl(T v, int src_lane) {
    static_assert(sizeof(T) <= 8, "shfl payload");
    Warp *w = tls.warp;
    unsigned long long bits = 0;
    memcpy(&bits, &v, sizeof(T));
    w->slot[tls.lane] = bits;
    w->bar.wait();
    T out = v;
    if (src_lane >= 0 && src_lane < w->bar.n) memcpy(&out, &w->slot[src_lane], sizeof(T));
    w->bar.wait();
    return out;
}

}  // namespace emu

#define threadIdx (emu::tls.threadIdx)
#define blockIdx (emu::tls.blockIdx)
#define blockDim (emu::tls.blockDim)
#define gridDim (emu::tls.gridDim)

// NOTE: an exited lane must not be waited for; kernels here only call collectives while all 32
// lanes of the warp are alive (true on the GPU side as well: full masks everywhere).
template <class T> static inline T __shfl_up_sync(unsigned, T v, int d) { return emu::shfl(v, emu::tls.lane - d); }
template <class T> static inline T __shfl_down_sync(unsigned, T v, int d) { return emu::shfl(v, emu::tls.lane + d); }
template <class T> static inline T __shfl_xor_sync(unsigned, T v, int m) { return emu::shfl(v, emu::tls.lane ^ m); }
template <class T> static inline T __shfl_sync(unsigned, T v, int l) { return emu::shfl(v, l); }
static inline bool __all_sync(unsigned, bool p) {
    // AND-reduce over the warp through the shuffle slots
    emu::Warp *w = emu::tls.warp;
    w->slot[emu::tls.lane] = p ? 1ull : 0ull;
    w->bar.wait();
    bool all = true;
    for (int i = 0; i < w->bar.n; ++i) all = all && w->slot[i];
    w->bar.wait();
    return all;
}
static inline bool __any_sync(unsigned m, bool p) { return !__all_sync(m, !p); }
static inline void __syncwarp(unsigned = 0xffffffffu) { emu::tls.warp->bar.wait(); }
static inline void __syncthreads() { emu::tls.cta->bar.wait(); }
static inline void __threadfence() { std::atomic_thread_fence(std::memory_order_seq_cst); }
static inline void __threadfence_system() { std::atomic_thread_fence(std::memory_order_seq_cst); }
static inline void __threadfence_block() { std::atomic_thread_fence(std::memory_order_seq_cst); }

template <class T> static inline T __ldg(const T *p) { return *p; }
static inline double rsqrt(double x) { return 1.0 / std::sqrt(x); }
static inline long long __double_as_longlong(double x) { long long b; memcpy(&b, &x, 8); return b; }
static inline double __longlong_as_double(long long b) { double x; memcpy(&x, &b, 8); return x; }

static inline unsigned atomicInc(unsigned *p, unsigned lim) {
    unsigned old = __atomic_load_n(p, __ATOMIC_SEQ_CST);
    for (;;) {
        const unsigned nv = (old >= lim) ? 0u : old + 1u;
        if (__atomic_compare_exchange_n(p, &old, nv, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) return old;
    }
}
static inline int atomicAdd(int *p, int v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
static inline unsigned atomicAdd(unsigned *p, unsigned v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
static inline unsigned atomicMax(unsigned *p, unsigned v) {
    unsigned old = __atomic_load_n(p, __ATOMIC_SEQ_CST);
    while (old < v && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
    return old;
}
static inline unsigned atomicOr(unsigned *p, unsigned v) { return __atomic_fetch_or(p, v, __ATOMIC_SEQ_CST); }
static inline int atomicExch(int *p, int v) { return __atomic_exchange_n(p, v, __ATOMIC_SEQ_CST); }
static inline double atomicAdd(double *p, double v) {
    unsigned long long *q = (unsigned long long *)p, old = __atomic_load_n(q, __ATOMIC_SEQ_CST);
    for (;;) {
        double cur; memcpy(&cur, &old, 8); cur += v;
        unsigned long long nv; memcpy(&nv, &cur, 8);
        if (__atomic_compare_exchange_n(q, &old, nv, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) { memcpy(&cur, &old, 8); return cur; }
    }
}

// __shared__ variables: one arena per declaration site and CTA.
#define BFD_SHARED(type, name) type &name = *static_cast<type *>(emu::tls.cta->get_shared(__COUNTER__, sizeof(type)))
#define BFD_SHARED_ARRAY(type, name, ...) \
    typedef type name##_t __VA_ARGS__;    \
    name##_t &name = *static_cast<name##_t *>(emu::tls.cta->get_shared(__COUNTER__, sizeof(name##_t)))

// kernel<<<grid, block, smem, stream>>>(args...)  ->  BFD_LAUNCH / BFD_LAUNCH_CORESIDENT
#define BFD_LAUNCH(kernel, grid, block, smem, stream, ...) \
    emu::launch(dim3(grid), dim3(block), false, [&] { kernel(__VA_ARGS__); })
#define BFD_LAUNCH_CORESIDENT(kernel, grid, block, smem, stream, ...) \
    emu::launch(dim3(grid), dim3(block), true, [&] { kernel(__VA_ARGS__); })
