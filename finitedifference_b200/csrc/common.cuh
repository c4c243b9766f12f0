// Shared helpers for the sm_100a Burgers kernels.
#pragma once
#include <stdint.h>
#include <stdio.h>

#ifdef BFD_EMULATE
// CPU SIMT emulation of these same sources, used only by the CPU test-suite (tests/simt_emu).
#include "cuda_emu.h"
#define BFD_GRID(n) ((n) < 2 ? (n) : 2)
#else
#include <cuda_runtime.h>
#define BFD_SHARED(type, name) __shared__ type name
#define BFD_SHARED_ARRAY(type, name, ...) __shared__ __align__(16) type name __VA_ARGS__
#define BFD_LAUNCH(kernel, grid, block, smem, stream, ...) kernel<<<grid, block, smem, stream>>>(__VA_ARGS__)
// kernels whose CTAs wait on one another: every CTA must be resident or ordered by ticket
#define BFD_LAUNCH_CORESIDENT(kernel, grid, block, smem, stream, ...) \
    kernel<<<grid, block, smem, stream>>>(__VA_ARGS__)
#define BFD_GRID(n) (n)
#endif

#include "../../include/burgers_b200.h"

namespace bfd {

extern uint64_t g_launches;     // kernels launched by this library (bench.py: gpu_launches)
extern char g_cuda_error[256];  // last CUDA error text

int record_cuda_error(cudaError_t e, const char *where);

#define BFD_CUDA(expr)                                                    \
    do {                                                                  \
        cudaError_t _e = (expr);                                          \
        if (_e != cudaSuccess) return bfd::record_cuda_error(_e, #expr); \
    } while (0)

#define BFD_LAUNCHED()                                                          \
    do {                                                                        \
        __atomic_fetch_add(&bfd::g_launches, 1, __ATOMIC_RELAXED);                                                      \
        cudaError_t _e = cudaGetLastError();                                    \
        if (_e != cudaSuccess) return bfd::record_cuda_error(_e, "kernel launch"); \
    } while (0)

constexpr int kSkew = 31;  // lane l of a wavefront warp runs l rows behind lane 0

// ------------------------------------------------------------------------------------------
// Skewed strip layout (SSL) of one state (u, v) in HBM.
//
// The grid is cut into vertical strips of W = 32*CPL columns.  A warp owns a strip and marches
// down the rows; lane l owns columns [l*CPL, (l+1)*CPL) of the strip and is l rows behind lane 0
// (the left neighbour's row must be finished first).  At step s the warp therefore touches the
// anti-diagonal  { (r = s - l, c = strip*W + l*CPL + q) }.  The SSL stores exactly that set
// contiguously ("slot-row" s), so every load and store of the wavefront is one fully coalesced
// 256*CPL-byte access per component -- no shared-memory transposition, no strided sectors:
//
//   element (comp, r, c):  strip = c / W, l = (c % W) / CPL, q = c % CPL, s = r + l
//   offset = ((strip*slot_rows + s)*2 + comp)*W + l*CPL + q          [doubles]
//
// Entries whose (r, c) falls outside the grid are padding and are kept at zero.
// ------------------------------------------------------------------------------------------
struct Ssl {
    int nx, ny, cpl, W, nstrips, slot_rows;
    size_t strip_stride;  // doubles per strip
    size_t state_doubles; // doubles per state (all strips)
};

__host__ __device__ inline Ssl make_ssl(int nx, int ny, int cpl) {
    Ssl L;
    L.nx = nx; L.ny = ny; L.cpl = cpl; L.W = 32 * cpl;
    L.nstrips = (nx + L.W - 1) / L.W;
    L.slot_rows = ((ny + kSkew + 7) / 8) * 8;
    L.strip_stride = (size_t)L.slot_rows * 2 * L.W;
    L.state_doubles = L.strip_stride * L.nstrips;
    return L;
}

__host__ __device__ inline size_t ssl_offset(const Ssl &L, int comp, int r, int c) {
    const int strip = c / L.W, cc = c % L.W, l = cc / L.cpl;
    return ((size_t)(strip * L.slot_rows + r + l) * 2 + comp) * L.W + cc;
}

// Quiet-NaN with an unusual payload: "not written yet".  Arithmetic NaNs are canonical
// (0x7ff8000000000000 / 0xfff8...), so a computed value never aliases it.
constexpr unsigned long long kSentinelBits = 0x7ff8dead0b200b20ULL;

}  // namespace bfd
