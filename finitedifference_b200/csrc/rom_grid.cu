// J(w) V on the WHOLE structured grid (the LSPG PROM and the ECSW training matrix): the same rows
// as rom_jv_kernel (csrc/rom.cu), computed by walking the grid instead of gathering through index
// tables.
//
// Replaces inviscid_burgers_exact_jac2D (BurgersFD_CleanFine/hypernet2D.py:2703-2732) followed by
// J.dot(basis) at :1993 (gauss_newton_LSPG) and :2812 (compute_ECSW_training_matrix_2D).
//
// Why a second kernel: the gather version reads 8 basis rows per (cell, mode) from L2 -- own,
// left, and the two cells of the y-difference, u and v each -- i.e. 4x the basis per call, and
// runs at the L2's pace (0.91 ms at 750^2, k = 95: 29 % of the HBM roof).  On the full grid the
// neighbours are known:
//   * left (r, c-1): the cell this warp handled one step earlier -> registers;
//   * y-difference cells A(r, c), B(r, c): B(r, c) = A(r-1, c) both for the exact operator
//     (A = own cell, get_ops :2509-2520) and for the operator inviscid_burgers_implicit2D_LSPG
//     builds at :167 (A = the transposed cell (c, r)), so a warp that owns R
//     consecutive grid rows loads R+1 A-rows for R rows of output.
// A warp walks a segment of columns for R rows; lane = mode (MC chunks of 32 modes per lane).
// Loads per (cell, mode): 2 (R+1)/R for the exact operator (3 at R = 2), 2 + 2 (R+1)/R for the
// :167 operator (5 at R = 2) instead of 8.  Measured at 750^2, k = 95 (profiles/r01_bench_rom.jsonl):
// 0.44 ms (exact, was 0.89) and 0.58 ms (:167, was 0.91).  HBM: V once (+ once more along the transposed walk
// for :167) + J V written.
#include "wavefront.cuh"
#include <stdlib.h>

namespace bfd {

struct JvGridArgs {
    int nx, ny, k, seg_len, nseg, ngroups;
    const double *cx, *cxl, *cy, *cyd;      // per row i = r*nx + c (bfd_rom_mesh tables)
    const double *w, *V, *rw;               // state, basis [2 nP][k] on the P cells, optional residual for the C block
    double *JV;                             // [2 n_out][k]
    // Slab view (whole grid: r_lo = 0, r_hi = ny, p_row0 = a_row0 = 0, a_len = ny, wA = w, VA = V).
    // Owned grid rows [r_lo, r_hi): outputs and coefficient tables are indexed (r - r_lo)*nx + c.
    // P arrays (w, V) hold grid rows [p_row0, r_hi) in row order: cell (r, c) at (r - p_row0)*nx + c,
    // v component nP cells further.  A arrays (wA, VA; the :167 operator only) hold the transposed
    // cells as a column slab: cell (c, r), r in [a_row0, r_hi), at c*a_len + (r - a_row0), v at + nA.
    int r_lo, r_hi, p_row0, a_row0, a_len;
    size_t nP, nA, n_out;
    const double *wA, *VA;
    double *C;                              // CBLOCK: one snapshot's k x N block of the training matrix
    size_t ldc;
};

constexpr int kCSeg = 8;                    // columns per segment of the CBLOCK variant (one 64-byte run per mode)

// CBLOCK (R = 1, exact operator): instead of storing the J V rows, form
//   C[j, cell] = r_u[cell] (J V)_u[cell, j] + r_v[cell] (J V)_v[cell, j]      (:2813-2814)
// for the kCSeg cells of the segment, transpose the kCSeg x 32 MC tile through shared memory and
// write it along cells -- compute_ECSW_training_matrix_2D's cell loop without materialising J V.
template <int R, int MC, bool TRANSPOSED, bool CBLOCK = false>
__global__ void __launch_bounds__(128) rom_jv_grid_kernel(const JvGridArgs a) {
    BFD_SHARED_ARRAY(double, ctile, [CBLOCK ? 4 : 1][CBLOCK ? kCSeg : 1][CBLOCK ? 32 * MC + 1 : 1]);
    const int nx = a.nx, k = a.k, r_hi = a.r_hi;
    const size_t nP = a.nP, nA = TRANSPOSED ? a.nA : a.nP, n_out = a.n_out;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int jbase = blockIdx.y * (32 * MC) + lane;
    bool jok[MC];
    int joff[MC];
#pragma unroll
    for (int m = 0; m < MC; ++m) {
        jok[m] = jbase + 32 * m < k;
        joff[m] = jok[m] ? jbase + 32 * m : 0;
    }
    const double *__restrict__ V = a.V;
    const double *__restrict__ w = a.w;
    const double *__restrict__ VA = TRANSPOSED ? a.VA : a.V;
    const double *__restrict__ wA = TRANSPOSED ? a.wA : a.w;
    const long long ntasks = (long long)a.ngroups * a.nseg;
    for (long long task = (long long)blockIdx.x * 4 + warp; task < ntasks; task += (long long)gridDim.x * 4) {
        const int g = (int)(task / a.nseg), s = (int)(task % a.nseg);
        const int r0 = a.r_lo + g * R, c0 = s * a.seg_len;
        const int c1 = c0 + a.seg_len < nx ? c0 + a.seg_len : nx;
        // left neighbour (own cell of the previous column), per owned row
        double Lu[R][MC], Lv[R][MC], wLu[R], wLv[R];
#pragma unroll
        for (int rr = 0; rr < R; ++rr) {
            const int r = r0 + rr;
            const bool ok = c0 > 0 && r < r_hi;
            const size_t cell = ok ? (size_t)(r - a.p_row0) * nx + (c0 - 1) : 0;
            wLu[rr] = ok ? w[cell] : 0.0;
            wLv[rr] = ok ? w[nP + cell] : 0.0;
#pragma unroll
            for (int m = 0; m < MC; ++m) {
                Lu[rr][m] = ok ? V[cell * k + joff[m]] : 0.0;
                Lv[rr][m] = ok ? V[(nP + cell) * k + joff[m]] : 0.0;
            }
        }
        // One column's operands: A rows r0-1 .. r0+R-1 of the y-difference and (the :167 operator
        // only) the own rows.  The next column is loaded at the END of a step, right after the stores
        // (measured: the :167 operator 0.49 -> 0.44 ms against loading at the top of the step; loading
        // it BEFORE the arithmetic of the current column -- a register prefetch -- was slower: 0.53 ms,
        // 158-242 registers).
        struct Col {
            double Au[R + 1][MC], Av[R + 1][MC], wAu[R + 1], wAv[R + 1];
            double Pu[TRANSPOSED ? R : 1][MC], Pv[TRANSPOSED ? R : 1][MC], wPu[TRANSPOSED ? R : 1], wPv[TRANSPOSED ? R : 1];
        };
        auto load_col = [&](int c, Col &d) {
#pragma unroll
            for (int rr = 0; rr <= R; ++rr) {
                const int r = r0 - 1 + rr;
                const bool ok = r >= 0 && r < r_hi;
                const size_t cell = !ok ? 0 : TRANSPOSED ? (size_t)c * a.a_len + (r - a.a_row0)
                                                         : (size_t)(r - a.p_row0) * nx + c;
                d.wAu[rr] = ok ? wA[cell] : 0.0;
                d.wAv[rr] = ok ? wA[nA + cell] : 0.0;
#pragma unroll
                for (int m = 0; m < MC; ++m) {
                    d.Au[rr][m] = ok ? VA[cell * k + joff[m]] : 0.0;
                    d.Av[rr][m] = ok ? VA[(nA + cell) * k + joff[m]] : 0.0;
                }
            }
            if (TRANSPOSED) {
#pragma unroll
                for (int rr = 0; rr < R; ++rr) {
                    const int r = r0 + rr;
                    const bool ok = r < r_hi;
                    const size_t pc = ok ? (size_t)(r - a.p_row0) * nx + c : 0;
                    d.wPu[rr] = ok ? w[pc] : 0.0;
                    d.wPv[rr] = ok ? w[nP + pc] : 0.0;
#pragma unroll
                    for (int m = 0; m < MC; ++m) {
                        d.Pu[rr][m] = ok ? V[pc * k + joff[m]] : 0.0;
                        d.Pv[rr][m] = ok ? V[(nP + pc) * k + joff[m]] : 0.0;
                    }
                }
            }
        };
        Col cur;
        load_col(c0, cur);
        for (int c = c0; c < c1; ++c) {
            const bool more = c + 1 < c1;
            const auto &Au = cur.Au;
            const auto &Av = cur.Av;
            const auto &wAu = cur.wAu;
            const auto &wAv = cur.wAv;
#pragma unroll
            for (int rr = 0; rr < R; ++rr) {
                const int r = r0 + rr;
                if (r >= r_hi) break;
                const size_t cell = (size_t)(r - a.r_lo) * nx + c;            // output / coefficient index
                double Pu[MC], Pv[MC], wPu, wPv;
                if (TRANSPOSED) {
                    wPu = cur.wPu[rr];
                    wPv = cur.wPv[rr];
#pragma unroll
                    for (int m = 0; m < MC; ++m) { Pu[m] = cur.Pu[rr][m]; Pv[m] = cur.Pv[rr][m]; }
                } else {
                    wPu = wAu[rr + 1];
                    wPv = wAv[rr + 1];
#pragma unroll
                    for (int m = 0; m < MC; ++m) { Pu[m] = Au[rr + 1][m]; Pv[m] = Av[rr + 1][m]; }
                }
                const double cx = a.cx[cell], cxl = c > 0 ? a.cxl[cell] : 0.0, cy = a.cy[cell],
                             cyd = r > 0 ? a.cyd[cell] : 0.0;
                // 2x2 blocks of the exact Jacobian (:2719-2732), same grouping as rom_jv_kernel
                const double pxu = cx * wPu, pxv = cx * wPv, lxu = cxl * wLu[rr], lxv = cxl * wLv[rr];
                const double ayu = cy * wAu[rr + 1], ayv = cy * wAv[rr + 1], byu = cyd * wAu[rr], byv = cyd * wAv[rr];
                double *ou = a.JV + cell * k, *ov = a.JV + (n_out + cell) * k;
#pragma unroll
                for (int m = 0; m < MC; ++m) {
                    const double vpu = Pu[m], vpv = Pv[m], vlu = Lu[rr][m], vlv = Lv[rr][m];
                    const double vau = Au[rr + 1][m], vav = Av[rr + 1][m], vbu = Au[rr][m], vbv = Av[rr][m];
                    const double ju = vpu + (pxu * vpu - lxu * vlu) + 0.5 * (ayv * vau - byv * vbu) + 0.5 * (ayu * vav - byu * vbv);
                    const double jv = vpv + 0.5 * (pxv * vpu - lxv * vlu) + (ayv * vav - byv * vbv) + 0.5 * (pxu * vpv - lxu * vlv);
                    if (CBLOCK) {
                        ctile[warp][c - c0][32 * m + lane] = a.rw[cell] * ju + a.rw[n_out + cell] * jv;
                    } else if (jok[m]) {
                        ou[joff[m]] = ju;
                        ov[joff[m]] = jv;
                    }
                    Lu[rr][m] = vpu;
                    Lv[rr][m] = vpv;
                }
                wLu[rr] = wPu;
                wLv[rr] = wPv;
            }
            if (more) load_col(c + 1, cur);
        }
        if (CBLOCK) {
            __syncwarp();
            const int ncol = c1 - c0, j0 = blockIdx.y * (32 * MC);
            for (int e = lane; e < kCSeg * 32 * MC; e += 32) {
                const int mode = e / kCSeg, cl = e % kCSeg;
                if (cl < ncol && j0 + mode < k && r0 < r_hi)
                    a.C[(size_t)(j0 + mode) * a.ldc + (size_t)(r0 - a.r_lo) * nx + c0 + cl] = ctile[warp][cl][mode];
            }
            __syncwarp();
        }
    }
}

template <int R, bool T>
static int launch_mc(const JvGridArgs &a, int mc, dim3 grid, cudaStream_t st) {
    switch (mc) {
    case 1: BFD_LAUNCH((rom_jv_grid_kernel<R, 1, T>), grid, 128, 0, st, a); break;
    case 2: BFD_LAUNCH((rom_jv_grid_kernel<R, 2, T>), grid, 128, 0, st, a); break;
    default: BFD_LAUNCH((rom_jv_grid_kernel<R, 3, T>), grid, 128, 0, st, a); break;
    }
    BFD_LAUNCHED();
    return BFD_OK;
}

static int launch_cblock(const JvGridArgs &a, int mc, dim3 grid, cudaStream_t st) {
    switch (mc) {
    case 1: BFD_LAUNCH((rom_jv_grid_kernel<1, 1, false, true>), grid, 128, 0, st, a); break;
    case 2: BFD_LAUNCH((rom_jv_grid_kernel<1, 2, false, true>), grid, 128, 0, st, a); break;
    default: BFD_LAUNCH((rom_jv_grid_kernel<1, 3, false, true>), grid, 128, 0, st, a); break;
    }
    BFD_LAUNCHED();
    return BFD_OK;
}

}  // namespace bfd

using namespace bfd;

static int jv_grid_launch(const bfd_rom_mesh *m, int nx, int ny, int transposed, int rows_per_warp, int r_lo, int r_hi,
                          const double *w, const double *V, const double *wA, const double *VA, int k, double *JV,
                          void *stream);

extern "C" int bfd_rom_jv_grid(const bfd_rom_mesh *m, int nx, int ny, int transposed, int rows_per_warp,
                               const double *w, const double *V, int k, double *JV, void *stream) {
    if (!m || nx < 1 || ny < 1 || m->nrows != (int64_t)nx * ny || m->ncells != m->nrows) return BFD_EINVAL;
    return jv_grid_launch(m, nx, ny, transposed, rows_per_warp, 0, ny, w, V, w, V, k, JV, stream);
}

extern "C" int bfd_rom_jv_grid_slab(const bfd_rom_mesh *m, int nx, int ny, int transposed, int r_lo, int r_hi,
                                    const double *wP, const double *VP, const double *wA, const double *VA, int k,
                                    double *JV, void *stream) {
    if (!m || nx < 1 || ny < 1 || r_lo < 0 || r_hi > ny || r_lo >= r_hi) return BFD_EINVAL;
    if (m->nrows != (int64_t)nx * (r_hi - r_lo)) return BFD_EINVAL;
    if (transposed && (!wA || !VA)) return BFD_EINVAL;
    return jv_grid_launch(m, nx, ny, transposed, 0, r_lo, r_hi, wP, VP, wA, VA, k, JV, stream);
}

static int jv_grid_launch(const bfd_rom_mesh *m, int nx, int ny, int transposed, int rows_per_warp, int r_lo, int r_hi,
                          const double *w, const double *V, const double *wA, const double *VA, int k, double *JV,
                          void *stream) {
    if (!m->cx || !m->cxl || !m->cy || !m->cyd || !w || !V || !JV || k < 1) return BFD_EINVAL;
    if (m->weight) return BFD_EINVAL;                       // weighted rows belong to the sampled mesh
    if (transposed && nx != ny) return BFD_EINVAL;          // :167 permutes a square grid
    const int R = rows_per_warp == 4 ? 4 : rows_per_warp == 2 ? 2 : 1;   // measured at 750^2, k = 95: 1 row per warp is fastest for both operators (0.34 / 0.44 ms), 4 rows slowest (registers)
    const int mc = k > 64 ? 3 : k > 32 ? 2 : 1;
    JvGridArgs a{};
    a.nx = nx; a.ny = ny; a.k = k;
    a.r_lo = r_lo; a.r_hi = r_hi;
    a.p_row0 = a.a_row0 = r_lo > 0 ? r_lo - 1 : 0;          // the slab arrays start one row below the owned rows
    a.a_len = r_hi - a.a_row0;
    a.nP = (size_t)nx * (r_hi - a.p_row0);
    a.nA = (size_t)nx * a.a_len;
    a.n_out = (size_t)nx * (r_hi - r_lo);
    a.wA = wA; a.VA = VA;
    a.ngroups = (r_hi - r_lo + R - 1) / R;
    // Short segments: the warps in flight then read neighbouring stretches of the basis, which is
    // what DRAM pages like (measured at 750^2, k = 95: segments of 75 columns 0.44 / 0.59 ms, of
    // 8-16 columns 0.34 / 0.49 ms for the exact / :167 operator); the price is one extra left
    // neighbour load from L2 per segment.
    a.seg_len = transposed ? 16 : 8;
    if (a.seg_len > nx) a.seg_len = nx;
    a.nseg = (nx + a.seg_len - 1) / a.seg_len;
    a.cx = m->cx; a.cxl = m->cxl; a.cy = m->cy; a.cyd = m->cyd;
    a.w = w; a.V = V; a.rw = nullptr; a.JV = JV;
    const long long ntasks = (long long)a.ngroups * a.nseg;
    dim3 grid((unsigned)BFD_GRID((int)((ntasks + 3) / 4)), (unsigned)((k + 32 * mc - 1) / (32 * mc)));
    cudaStream_t st = (cudaStream_t)stream;
    if (R == 1) return transposed ? launch_mc<1, true>(a, mc, grid, st) : launch_mc<1, false>(a, mc, grid, st);
    if (R == 2) return transposed ? launch_mc<2, true>(a, mc, grid, st) : launch_mc<2, false>(a, mc, grid, st);
    return transposed ? launch_mc<4, true>(a, mc, grid, st) : launch_mc<4, false>(a, mc, grid, st);
}

extern "C" int bfd_ecsw_c_grid(const bfd_rom_mesh *m, int nx, int ny, const double *w, const double *r,
                               const double *V, int k, double *Cblk, size_t ldc, void *stream) {
    if (!m || nx < 1 || ny < 1 || m->nrows != (int64_t)nx * ny || m->ncells != m->nrows) return BFD_EINVAL;
    if (!m->cx || !m->cxl || !m->cy || !m->cyd || !w || !r || !V || !Cblk || k < 1 || m->weight) return BFD_EINVAL;
    if (ldc < (size_t)m->nrows) return BFD_EINVAL;
    const int mc = k > 64 ? 3 : k > 32 ? 2 : 1;
    JvGridArgs a{};
    a.nx = nx; a.ny = ny; a.k = k;
    a.ngroups = ny;
    a.seg_len = kCSeg < nx ? kCSeg : nx;
    a.nseg = (nx + a.seg_len - 1) / a.seg_len;
    a.cx = m->cx; a.cxl = m->cxl; a.cy = m->cy; a.cyd = m->cyd;
    a.w = w; a.V = V; a.rw = r; a.JV = nullptr; a.C = Cblk; a.ldc = ldc;
    a.r_lo = 0; a.r_hi = ny; a.p_row0 = a.a_row0 = 0; a.a_len = ny;
    a.nP = a.nA = a.n_out = (size_t)nx * ny;
    a.wA = w; a.VA = V;
    const long long ntasks = (long long)a.ngroups * a.nseg;
    long long blocks = (ntasks + 3) / 4;
    if (blocks > 148 * 64) blocks = 148 * 64;
    dim3 grid((unsigned)BFD_GRID((int)blocks), (unsigned)((k + 32 * mc - 1) / (32 * mc)));
    return launch_cblock(a, mc, grid, (cudaStream_t)stream);
}
