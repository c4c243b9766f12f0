// Natural-layout operators: residual of one trapezoidal step on the (u | v) row-major state.
//
// Replaces inviscid_burgers_res2D (hypernet2D.py:2544-2586) and inviscid_burgers_res2D_alt
// (:2588-2646): the reference forms four sparse products and ~15 NumPy temporaries; here the
// stencil is evaluated per cell, one thread per cell along x so that u, v, up, vp and both
// outputs are coalesced; the left neighbour comes from the adjacent lane's registers (shuffle),
// the lower-row neighbour from L2.  HBM-bound: 32 B read + 16 B written per cell.
#include "common.cuh"

namespace bfd {

struct ResidualArgs {
    int nx, ny;
    const double *cx, *cxl, *cy, *cyd, *src, *lbc, *w, *wp;
    double *r, *norm2;
};

__global__ void __launch_bounds__(256) residual_kernel(const ResidualArgs a) {
    const size_t N = (size_t)a.nx * a.ny;
    const int lane = threadIdx.x & 31;
    double acc = 0.0;
    // one warp per 32-column segment of a row; grid-stride over segments
    const int segs_x = (a.nx + 31) / 32;
    const size_t nseg = (size_t)segs_x * a.ny;
    const size_t warp0 = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t seg = warp0; seg < nseg; seg += nwarp) {
        const int row = (int)(seg / segs_x);
        const int c = (int)(seg - (size_t)row * segs_x) * 32 + lane;
        const bool ok = c < a.nx;
        const int cc = ok ? c : a.nx - 1;
        const size_t p = (size_t)row * a.nx + cc;
        const double u = a.w[p], v = a.w[N + p], up = a.wp[p], vp = a.wp[N + p];
        const double F = 0.5 * (u * u) + 0.5 * (up * up);      // :2619
        const double H = 0.5 * (v * v) + 0.5 * (vp * vp);      // :2620
        const double G = 0.5 * u * v + 0.5 * up * vp;          // :2621
        // left neighbour: previous lane, or a reload at the segment boundary
        double FL = __shfl_up_sync(0xffffffffu, F, 1), GL = __shfl_up_sync(0xffffffffu, G, 1);
        if (lane == 0) {
            FL = GL = 0.0;
            if (c > 0 && ok) {
                const double ul = a.w[p - 1], vl = a.w[N + p - 1], upl = a.wp[p - 1], vpl = a.wp[N + p - 1];
                FL = 0.5 * (ul * ul) + 0.5 * (upl * upl);
                GL = 0.5 * ul * vl + 0.5 * upl * vpl;
            }
        }
        double GD = 0.0, HD = 0.0;
        if (row > 0) {
            const size_t d = p - a.nx;
            const double ud = a.w[d], vd = a.w[N + d], upd = a.wp[d], vpd = a.wp[N + d];
            GD = 0.5 * ud * vd + 0.5 * upd * vpd;
            HD = 0.5 * (vd * vd) + 0.5 * (vpd * vpd);
        }
        const double cx = a.cx[cc], cxl = a.cxl[cc], cy = a.cy[row], cyd = a.cyd[row];
        double ru = (u - up) + (cx * F - cxl * FL) + (cy * G - cyd * GD) - a.src[cc];   // :2637-2638
        if (c == 0) ru -= a.lbc[row];                                                   // :2639
        const double rv = (v - vp) + (cy * H - cyd * HD) + (cx * G - cxl * GL);         // :2642-2643
        if (ok) {
            a.r[p] = ru;
            a.r[N + p] = rv;
            acc += ru * ru + rv * rv;
        }
    }
    if (a.norm2) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
        if (lane == 0) atomicAdd(a.norm2, acc);
    }
}

}  // namespace bfd

using namespace bfd;

extern "C" int bfd_residual(int nx, int ny, const double *cx, const double *cxl, const double *cy,
                            const double *cyd, const double *src, const double *lbc, const double *w,
                            const double *wp, double *r, double *norm2, void *stream) {
    if (nx < 1 || ny < 1 || !cx || !cxl || !cy || !cyd || !src || !lbc || !w || !wp || !r) return BFD_EINVAL;
    ResidualArgs a{nx, ny, cx, cxl, cy, cyd, src, lbc, w, wp, r, norm2};
    const size_t nseg = (size_t)((nx + 31) / 32) * ny;
    size_t grid = (nseg + 7) / 8;
    if (grid > 148 * 8) grid = 148 * 8;
    BFD_LAUNCH(residual_kernel, BFD_GRID((int)grid), 256, 0, (cudaStream_t)stream, a);
    BFD_LAUNCHED();
    return BFD_OK;
}

// ------------------------------------------------------------------------------------------
// FP64 pipe probe: the DFMA issue rate of THIS device, measured in the run that reports a fraction
// of it (bench.py's roofline of the FOM sweep kernel, whose binding unit is the FP64 pipe).  Eight
// independent chains per thread, 32 warps per SM.  The caller times the launch with events;
// flop = 2 * 8 * iters * threads.  Measurement aid: no counterpart in the reference.
// ------------------------------------------------------------------------------------------
namespace bfd {
__global__ void __launch_bounds__(1024) dfma_probe_kernel(double *sink, int iters, double a, double b) {
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
#pragma unroll 4
    for (int i = 0; i < iters; ++i) {
        x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
        x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
    }
    sink[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}
}  // namespace bfd

extern "C" int bfd_probe_dfma(double *sink, int blocks, int iters, void *stream) {
    if (!sink || blocks < 1 || iters < 1) return BFD_EINVAL;
    BFD_LAUNCH(dfma_probe_kernel, BFD_GRID(blocks), 1024, 0, (cudaStream_t)stream, sink, iters, 0.999999, 1e-6);
    BFD_LAUNCHED();
    return BFD_OK;
}
