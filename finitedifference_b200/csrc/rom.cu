// Reduced-order models: gather-based residual and matrix-free Jacobian x basis product on an
// arbitrary set of mesh rows -- the whole grid for the LSPG PROM, the sampled cells for the
// ECSW/ECM HPROM.
//
// Replaces, for the reference's PROM / HPROM paths (BurgersFD_CleanFine/hypernet2D.py):
//   inviscid_burgers_res2D :2544-2586 and inviscid_burgers_res2D_ecsw :2648-2701  (residual rows),
//   inviscid_burgers_exact_jac2D :2703-2732 / _exact_jac2D_ecsw :2734-2744 followed by the
//   sparse x dense products J.dot(basis) at :1993, :2030 and :2812                (J V rows),
//   the cell loop of compute_ECSW_training_matrix_2D :2807-2814                   (C blocks).
// The reference assembles CSR matrices and multiplies them with the dense basis; here a row of
// J V is formed directly from the 2x2 blocks of the exact Jacobian and the basis rows of the
// (at most) four cells it couples to.  Which cells those are is data (index tables built by the
// host, finitedifference_b200/rom.py), so one kernel covers
//   - the exact operator (get_ops, :2509-2520): own cell, left cell, lower-row cell;
//   - the operator inviscid_burgers_implicit2D_LSPG really builds (:167, rows-only permutation):
//     its y-difference reads the TRANSPOSED cells (c, r) and (c, r-1)  [SURVEY.md Q1];
//   - the sampled mesh, where indices point into the augmented-cell arrays (:242-245).
// State layout: u of state cell q at w[q], v at w[ncells + q]; basis row-major [2*ncells][k].
// HBM-bound gathers: 8 basis-row reads + 2 row writes per (row, mode).
#include "common.cuh"

namespace bfd {

struct RomArgs {
    bfd_rom_mesh m;
    const double *w, *wp, *V;
    double *r, *norm2, *JV;
    int k;
};

// r = residual rows [sample u | sample v], times the row weight when given (:2020, :2037)
__global__ void __launch_bounds__(256) rom_residual_kernel(const RomArgs a) {
    const bfd_rom_mesh &m = a.m;
    const long long n = m.nrows, nc = m.ncells;
    double acc = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int p = m.p[i], xl = m.xl[i], yd = m.yd[i];
        const double u = a.w[p], v = a.w[nc + p], up = a.wp[p], vp = a.wp[nc + p];
        const double F = 0.5 * (u * u) + 0.5 * (up * up);       // :2619 / :2687
        const double H = 0.5 * (v * v) + 0.5 * (vp * vp);
        const double G = 0.5 * u * v + 0.5 * up * vp;
        double FL = 0.0, GL = 0.0, GD = 0.0, HD = 0.0;
        if (xl >= 0) {
            const double ul = a.w[xl], vl = a.w[nc + xl], upl = a.wp[xl], vpl = a.wp[nc + xl];
            FL = 0.5 * (ul * ul) + 0.5 * (upl * upl);
            GL = 0.5 * ul * vl + 0.5 * upl * vpl;
        }
        if (yd >= 0) {
            const double ud = a.w[yd], vd = a.w[nc + yd], upd = a.wp[yd], vpd = a.wp[nc + yd];
            GD = 0.5 * ud * vd + 0.5 * upd * vpd;
            HD = 0.5 * (vd * vd) + 0.5 * (vpd * vpd);
        }
        const double cx = m.cx[i], cxl = m.cxl[i], cy = m.cy[i], cyd = m.cyd[i];
        double ru = ((u - up) + (cx * F - cxl * FL) + (cy * G - cyd * GD) - m.src[i]) - m.lbc[i];
        double rv = (v - vp) + (cy * H - cyd * HD) + (cx * G - cxl * GL);
        if (m.weight) { const double wt = m.weight[i]; ru *= wt; rv *= wt; }
        a.r[i] = ru;
        a.r[n + i] = rv;
        acc += ru * ru + rv * rv;
    }
    if (a.norm2) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
        if ((threadIdx.x & 31) == 0) atomicAdd(a.norm2, acc);
    }
}

// (J V) rows [sample u | sample v] x k, times the row weight when given (:2030-2032).
// blockDim = (32, 8): x runs over modes (coalesced basis rows), y over rows of the tile.
__global__ void __launch_bounds__(256) rom_jv_kernel(const RomArgs a) {
    const bfd_rom_mesh &m = a.m;
    const long long n = m.nrows, nc = m.ncells;
    const int k = a.k;
    for (long long i = blockIdx.x * (long long)blockDim.y + threadIdx.y; i < n; i += (long long)gridDim.x * blockDim.y) {
        const int p = m.p[i], xl = m.xl[i], ja = m.ja[i], jb = m.jb[i];
        const double cx = m.cx[i], cxl = (xl >= 0) ? m.cxl[i] : 0.0, cy = (ja >= 0) ? m.cy[i] : 0.0,
                     cyd = (jb >= 0) ? m.cyd[i] : 0.0;
        const int xls = xl >= 0 ? xl : p, jas = ja >= 0 ? ja : p, jbs = jb >= 0 ? jb : p;
        const double u = a.w[p], v = a.w[nc + p];
        const double ul = a.w[xls], vl = a.w[nc + xls];
        const double ua = a.w[jas], va = a.w[nc + jas], ub = a.w[jbs], vb = a.w[nc + jbs];
        const double wt = m.weight ? m.weight[i] : 1.0;
        // 2x2 blocks of the exact Jacobian (:2719-2732) at the cells this row couples to
        const double pxu = cx * u, pxv = cx * v, lxu = cxl * ul, lxv = cxl * vl;     // x-difference: own, left
        const double ayu = cy * ua, ayv = cy * va, byu = cyd * ub, byv = cyd * vb;   // y-difference: its two cells
        const double *Vp = a.V + (size_t)p * k, *Vpv = a.V + (size_t)(nc + p) * k;
        const double *Vl = a.V + (size_t)xls * k, *Vlv = a.V + (size_t)(nc + xls) * k;
        const double *Va = a.V + (size_t)jas * k, *Vav = a.V + (size_t)(nc + jas) * k;
        const double *Vb = a.V + (size_t)jbs * k, *Vbv = a.V + (size_t)(nc + jbs) * k;
        double *ou = a.JV + (size_t)i * k, *ov = a.JV + (size_t)(n + i) * k;
        for (int j = threadIdx.x; j < k; j += 32) {
            const double vpu = Vp[j], vpv = Vpv[j], vlu = Vl[j], vlv = Vlv[j];
            const double vau = Va[j], vav = Vav[j], vbu = Vb[j], vbv = Vbv[j];
            // u rows: I + JDx diag(u) + 0.5 JDy diag(v) on the u columns, 0.5 JDy diag(u) on the v columns
            const double ju = vpu + (pxu * vpu - lxu * vlu) + 0.5 * (ayv * vau - byv * vbu) + 0.5 * (ayu * vav - byu * vbv);
            // v rows: 0.5 JDx diag(v) on the u columns, I + JDy diag(v) + 0.5 JDx diag(u) on the v columns
            const double jv = vpv + 0.5 * (pxv * vpu - lxv * vlu) + (ayv * vav - byv * vbv) + 0.5 * (pxu * vpv - lxu * vlv);
            ou[j] = wt * ju;
            ov[j] = wt * jv;
        }
    }
}

// The Jacobian itself, for callers that want the matrix (inviscid_burgers_exact_jac2D, :2703-2732):
// 12 values per row pair in fixed slots; the host owns the column pattern
//   u row:  (p) (xl) (ja) (jb) | (N+ja) (N+jb)      v row:  (p) (xl) | (N+p) (N+xl) (N+ja) (N+jb)
__global__ void __launch_bounds__(256) rom_jac_values_kernel(const RomArgs a) {
    const bfd_rom_mesh &m = a.m;
    const long long n = m.nrows, nc = m.ncells;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int p = m.p[i], xl = m.xl[i], ja = m.ja[i], jb = m.jb[i];
        const double cx = m.cx[i], cxl = (xl >= 0) ? m.cxl[i] : 0.0, cy = (ja >= 0) ? m.cy[i] : 0.0,
                     cyd = (jb >= 0) ? m.cyd[i] : 0.0;
        const int xls = xl >= 0 ? xl : p, jas = ja >= 0 ? ja : p, jbs = jb >= 0 ? jb : p;
        const double pxu = cx * a.w[p], pxv = cx * a.w[nc + p], lxu = cxl * a.w[xls], lxv = cxl * a.w[nc + xls];
        const double ayu = cy * a.w[jas], ayv = cy * a.w[nc + jas], byu = cyd * a.w[jbs], byv = cyd * a.w[nc + jbs];
        double *o = a.JV + 12 * i;
        o[0] = 1.0 + pxu;  o[1] = -lxu;       o[2] = 0.5 * ayv;        o[3] = -0.5 * byv;  o[4] = 0.5 * ayu;  o[5] = -0.5 * byu;
        o[6] = 0.5 * pxv;  o[7] = -0.5 * lxv; o[8] = 1.0 + 0.5 * pxu;  o[9] = -0.5 * lxu;  o[10] = ayv;       o[11] = -byv;
    }
}

// C[j, cell] = r_u[cell]*(JV)_u[cell, j] + r_v[cell]*(JV)_v[cell, j]   (:2813-2814), one snapshot's
// k x nrows block written row-major with leading dimension ldc; 32x32 tiles through shared memory
// so that reads run along modes and writes along cells.
__global__ void __launch_bounds__(256) ecsw_c_block_kernel(long long n, int k, const double *r, const double *JV,
                                                           double *C, size_t ldc) {
    BFD_SHARED_ARRAY(double, tile, [32][33]);
    const long long cell0 = (long long)blockIdx.x * 32;
    const int j0 = blockIdx.y * 32;
    for (int cy = threadIdx.y; cy < 32; cy += blockDim.y) {
        const long long cell = cell0 + cy;
        const int j = j0 + threadIdx.x;
        double val = 0.0;
        if (cell < n && j < k) val = r[cell] * JV[(size_t)cell * k + j] + r[n + cell] * JV[(size_t)(n + cell) * k + j];
        tile[cy][threadIdx.x] = val;
    }
    __syncthreads();
    for (int jy = threadIdx.y; jy < 32; jy += blockDim.y) {
        const int j = j0 + jy;
        const long long cell = cell0 + threadIdx.x;
        if (cell < n && j < k) C[(size_t)j * ldc + cell] = tile[threadIdx.x][jy];
    }
}

// The same block without the intermediate J V: a 32-cell x 32-mode tile of
// C[j, cell] = r_u (J V)_u[cell, j] + r_v (J V)_v[cell, j] is formed from the basis rows directly
// (reads along modes), transposed in shared memory and written along cells.  HBM traffic per
// snapshot: V once + the C block, instead of V + 3 x (J V) + C.
__global__ void __launch_bounds__(256) ecsw_c_fused_kernel(const RomArgs a, const double *r, double *C, size_t ldc) {
    BFD_SHARED_ARRAY(double, tile, [32][33]);
    const bfd_rom_mesh &m = a.m;
    const long long n = m.nrows, nc = m.ncells;
    const int k = a.k;
    const long long cell0 = (long long)blockIdx.x * 32;
    const int j = blockIdx.y * 32 + threadIdx.x;
    for (int cy = threadIdx.y; cy < 32; cy += blockDim.y) {
        const long long i = cell0 + cy;
        double val = 0.0;
        if (i < n && j < k) {
            const int p = m.p[i], xl = m.xl[i], ja = m.ja[i], jb = m.jb[i];
            const double cx = m.cx[i], cxl = (xl >= 0) ? m.cxl[i] : 0.0, cy_ = (ja >= 0) ? m.cy[i] : 0.0,
                         cyd = (jb >= 0) ? m.cyd[i] : 0.0;
            const int xls = xl >= 0 ? xl : p, jas = ja >= 0 ? ja : p, jbs = jb >= 0 ? jb : p;
            const double pxu = cx * a.w[p], pxv = cx * a.w[nc + p], lxu = cxl * a.w[xls], lxv = cxl * a.w[nc + xls];
            const double ayu = cy_ * a.w[jas], ayv = cy_ * a.w[nc + jas], byu = cyd * a.w[jbs], byv = cyd * a.w[nc + jbs];
            const double vpu = a.V[(size_t)p * k + j], vpv = a.V[(size_t)(nc + p) * k + j];
            const double vlu = a.V[(size_t)xls * k + j], vlv = a.V[(size_t)(nc + xls) * k + j];
            const double vau = a.V[(size_t)jas * k + j], vav = a.V[(size_t)(nc + jas) * k + j];
            const double vbu = a.V[(size_t)jbs * k + j], vbv = a.V[(size_t)(nc + jbs) * k + j];
            const double ju = vpu + (pxu * vpu - lxu * vlu) + 0.5 * (ayv * vau - byv * vbu) + 0.5 * (ayu * vav - byu * vbv);
            const double jv = vpv + 0.5 * (pxv * vpu - lxv * vlu) + (ayv * vav - byv * vbv) + 0.5 * (pxu * vpv - lxu * vlv);
            val = r[i] * ju + r[n + i] * jv;
        }
        tile[cy][threadIdx.x] = val;
    }
    __syncthreads();
    for (int jy = threadIdx.y; jy < 32; jy += blockDim.y) {
        const int jj = blockIdx.y * 32 + jy;
        const long long cell = cell0 + threadIdx.x;
        if (cell < n && jj < k) C[(size_t)jj * ldc + cell] = tile[threadIdx.x][jy];
    }
}

}  // namespace bfd

using namespace bfd;

static int check_mesh(const bfd_rom_mesh *m) {
    if (!m || m->nrows < 1 || m->ncells < 1 || !m->p || !m->xl || !m->yd || !m->ja || !m->jb) return BFD_EINVAL;
    if (!m->cx || !m->cxl || !m->cy || !m->cyd || !m->src || !m->lbc) return BFD_EINVAL;
    return BFD_OK;
}

extern "C" int bfd_rom_residual(const bfd_rom_mesh *m, const double *w, const double *wp, double *r, double *norm2,
                                void *stream) {
    if (check_mesh(m) != BFD_OK || !w || !wp || !r) return BFD_EINVAL;
    RomArgs a{*m, w, wp, nullptr, r, norm2, nullptr, 0};
    long long grid = (m->nrows + 255) / 256;
    if (grid > 148 * 8) grid = 148 * 8;
    BFD_LAUNCH(rom_residual_kernel, BFD_GRID((int)grid), 256, 0, (cudaStream_t)stream, a);
    BFD_LAUNCHED();
    return BFD_OK;
}

extern "C" int bfd_rom_jv(const bfd_rom_mesh *m, const double *w, const double *V, int k, double *JV, void *stream) {
    if (check_mesh(m) != BFD_OK || !w || !V || !JV || k < 1) return BFD_EINVAL;
    RomArgs a{*m, w, nullptr, V, nullptr, nullptr, JV, k};
    long long grid = (m->nrows + 7) / 8;
    if (grid > 148 * 16) grid = 148 * 16;
    BFD_LAUNCH(rom_jv_kernel, BFD_GRID((int)grid), dim3(32, 8), 0, (cudaStream_t)stream, a);
    BFD_LAUNCHED();
    return BFD_OK;
}

extern "C" int bfd_jac_csr_values(const bfd_rom_mesh *m, const double *w, double *vals, void *stream) {
    if (check_mesh(m) != BFD_OK || !w || !vals) return BFD_EINVAL;
    RomArgs a{*m, w, nullptr, nullptr, nullptr, nullptr, vals, 0};
    long long grid = (m->nrows + 255) / 256;
    if (grid > 148 * 8) grid = 148 * 8;
    BFD_LAUNCH(rom_jac_values_kernel, BFD_GRID((int)grid), 256, 0, (cudaStream_t)stream, a);
    BFD_LAUNCHED();
    return BFD_OK;
}

extern "C" int bfd_ecsw_c_fused(const bfd_rom_mesh *m, const double *w, const double *r, const double *V, int k,
                                double *C, size_t ldc, void *stream) {
    if (check_mesh(m) != BFD_OK || !w || !r || !V || !C || k < 1 || ldc < (size_t)m->nrows) return BFD_EINVAL;
    RomArgs a{*m, w, nullptr, V, nullptr, nullptr, nullptr, k};
    dim3 grid((unsigned)((m->nrows + 31) / 32), (unsigned)((k + 31) / 32)), block(32, 8);
    BFD_LAUNCH(ecsw_c_fused_kernel, grid, block, 0, (cudaStream_t)stream, a, r, C, ldc);
    BFD_LAUNCHED();
    return BFD_OK;
}

extern "C" int bfd_ecsw_c_block(int64_t nrows, int k, const double *r, const double *JV, double *C, size_t ldc,
                                void *stream) {
    if (nrows < 1 || k < 1 || !r || !JV || !C || ldc < (size_t)nrows) return BFD_EINVAL;
    dim3 grid((unsigned)((nrows + 31) / 32), (unsigned)((k + 31) / 32)), block(32, 8);
    BFD_LAUNCH(ecsw_c_block_kernel, grid, block, 0, (cudaStream_t)stream, (long long)nrows, k, r, JV, C, ldc);
    BFD_LAUNCHED();
    return BFD_OK;
}
