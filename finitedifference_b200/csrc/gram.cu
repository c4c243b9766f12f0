// Normal equations of the LSPG Gauss-Newton step on the FP64 tensor cores:
//     G = X^T X  (k x k),   b = X^T f  (k),   X = J V  (M x k row-major, M = 2N rows, k <= 96).
//
// Replaces numpy.linalg.lstsq(JV, -f) of gauss_newton_LSPG (BurgersFD_CleanFine/hypernet2D.py:1994):
// with cond(JV) ~ 1.5 the k x k normal equations give the same step to 1e-13, and
// forming them is the dense part of the PROM: 2 M k^2 flop (20.3 GFLOP at 750^2, k = 95).
// cuBLAS runs this tall-skinny shape (95 x 95 output, 1.1 M deep) at 22 TFLOP/s (0.92 ms); it cannot
// use the symmetry of G.  Here:
//   * X streams through shared memory in tiles of 16 rows (cp.async, 3 stages), padded to 96 modes
//     plus one extra column block that carries f, so b falls out of the same contraction;
//   * G is cut into 8 x 8 blocks; only the 78 blocks on or above the diagonal (+ 12 for b) are
//     computed, with DMMA (mma.sync.m8n8k4.f64: FP64 tensor pipe).  Warp I of a CTA owns block rows
//     I and 11-I (15 blocks each way: balanced), and the A fragment of block row i is the B
//     fragment of block column i, so a warp loads at most 13 fragments for 15 DMMAs per 4 rows;
//   * every CTA keeps its partial G in registers over all its tiles and writes it once; a second
//     small kernel adds the partials in a fixed order (deterministic: replicas and ranks agree
//     bit for bit) and mirrors the upper triangle.
#include "wavefront.cuh"
#include <stdlib.h>

namespace bfd {

constexpr int kGramRows = 16;        // rows of X per stage (4 DMMA k-steps)
constexpr int kGramPitch = 108;      // doubles per shared row: 96 modes + f + pad; 108 mod 16 = 12 -> fragment loads conflict-free
constexpr int kGramStages = 3;
constexpr int kGramWarps = 6;        // block rows (I, 11-I), I = 0..5
constexpr int kGramBlocks = 13;      // 12 mode blocks + the f block
constexpr int kGramLd = 104;         // leading dimension of a partial: 13 blocks x 8

struct GramArgs {
    const double *X, *f;
    long long M;
    int k, ntiles;
    double *partial;                 // [gridDim.x][96][kGramLd]
};

#ifndef BFD_EMULATE
__device__ __forceinline__ void dmma8x8x4(double (&d)[2], double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                 : "+d"(d[0]), "+d"(d[1]) : "d"(a), "d"(b));
}
#else
// D[m][n] += sum_k A[m][k] B[k][n]; lane l holds A[l/4][l%4], B[l%4][l/4], D[l/4][2*(l%4) + {0,1}]
inline void dmma8x8x4(double (&d)[2], double a, double b) {
    const int l = emu::tls.lane;
    for (int kk = 0; kk < 4; ++kk) {
        const double av = __shfl_sync(0xffffffffu, a, (l / 4) * 4 + kk);
        const double b0 = __shfl_sync(0xffffffffu, b, (2 * (l % 4)) * 4 + kk);
        const double b1 = __shfl_sync(0xffffffffu, b, (2 * (l % 4) + 1) * 4 + kk);
        d[0] += av * b0;
        d[1] += av * b1;
    }
}
#endif

// One warp's share of a 16-row tile: block rows I (columns I..12, acc[0..N1)) and 11-I (columns
// 11-I..12, acc[N1..15)).  The six instantiations run interleaved on every SM sub-partition, so
// the loop is kept short (no unrolling over the k-steps: with 4x unrolling the kernel spent 6
// cycles per issue waiting for instructions, profiles/r01_gram_ncu.md).
template <int I>
__device__ __forceinline__ void gram_compute(double (&acc)[15][2], saddr sb) {
    constexpr int I2 = 11 - I, N1 = kGramBlocks - I, N2 = kGramBlocks - I2;
#pragma unroll 1
    for (int ks = 0; ks < kGramRows / 4; ++ks) {
        double fr[N1];                                       // fragments of column blocks I..12
#pragma unroll
        for (int j = 0; j < N1; ++j) fr[j] = lds(sb + (saddr)((ks * 4 * kGramPitch + (I + j) * 8) * 8));
#pragma unroll
        for (int j = 0; j < N1; ++j) dmma8x8x4(acc[j], fr[0], fr[j]);
#pragma unroll
        for (int j = 0; j < N2; ++j) dmma8x8x4(acc[N1 + j], fr[I2 - I], fr[I2 - I + j]);
    }
}

// lane holds rows (l >> 2), columns 2 (l & 3) + {0, 1} of each of its blocks
template <int I>
__device__ __forceinline__ void gram_store(const double (&acc)[15][2], double *P, int lane) {
    constexpr int I2 = 11 - I, N1 = kGramBlocks - I, N2 = kGramBlocks - I2;
#pragma unroll
    for (int j = 0; j < N1; ++j) {
        double *p = P + (size_t)(8 * I + (lane >> 2)) * kGramLd + 8 * (I + j) + 2 * (lane & 3);
        p[0] = acc[j][0]; p[1] = acc[j][1];
    }
#pragma unroll
    for (int j = 0; j < N2; ++j) {
        double *p = P + (size_t)(8 * I2 + (lane >> 2)) * kGramLd + 8 * (I2 + j) + 2 * (lane & 3);
        p[0] = acc[N1 + j][0]; p[1] = acc[N1 + j][1];
    }
}

// stage `st` <- tile `nt` of X (and its slice of f in column 96); rows past M are zero
__device__ __forceinline__ void gram_load(const GramArgs &a, saddr tiles, int st, long long nt) {
    if (nt < a.ntiles) {
        const long long r0 = nt * kGramRows;
        const saddr sb = tiles + (saddr)(st * kGramRows * kGramPitch * 8);
        const int k = a.k;
        for (int e = threadIdx.x; e < kGramRows * k; e += 32 * kGramWarps) {
            const int row = e / k, col = e - row * k;
            const saddr dst = sb + (saddr)((row * kGramPitch + col) * 8);
            if (r0 + row < a.M) cp_async8(dst, a.X + r0 * k + e); else sts(dst, 0.0);
        }
        if (threadIdx.x < kGramRows) {
            const saddr dst = sb + (saddr)((threadIdx.x * kGramPitch + 96) * 8);
            if (r0 + threadIdx.x < a.M) cp_async8(dst, a.f + r0 + threadIdx.x); else sts(dst, 0.0);
        }
    }
    cp_async_commit();
}

__global__ void __launch_bounds__(32 * kGramWarps) gram_partial_kernel(const GramArgs a) {
    BFD_SHARED_ARRAY(double, xs, [kGramStages * kGramRows * kGramPitch]);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const saddr tiles = smem_addr(&xs[0]);
    for (int e = threadIdx.x; e < kGramStages * kGramRows * kGramPitch; e += 32 * kGramWarps) xs[e] = 0.0;   // pad columns stay 0
    __syncthreads();
    for (int s = 0; s < kGramStages - 1; ++s) gram_load(a, tiles, s, (long long)blockIdx.x + (long long)s * gridDim.x);
    double acc[15][2];
#pragma unroll
    for (int j = 0; j < 15; ++j) acc[j][0] = acc[j][1] = 0.0;
    const saddr lane_off = (saddr)(((lane & 3) * kGramPitch + (lane >> 2)) * 8);
    int it = 0;
    for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x, ++it) {
        cp_async_wait_pending<kGramStages - 2>();
        __syncthreads();
        // refill the stage consumed one tile ago with the tile kGramStages-1 ahead
        gram_load(a, tiles, (it + kGramStages - 1) % kGramStages, (long long)tile + (long long)(kGramStages - 1) * gridDim.x);
        const saddr sb = tiles + (saddr)((it % kGramStages) * kGramRows * kGramPitch * 8) + lane_off;
        switch (warp) {
        case 0: gram_compute<0>(acc, sb); break;
        case 1: gram_compute<1>(acc, sb); break;
        case 2: gram_compute<2>(acc, sb); break;
        case 3: gram_compute<3>(acc, sb); break;
        case 4: gram_compute<4>(acc, sb); break;
        default: gram_compute<5>(acc, sb); break;
        }
    }
    double *P = a.partial + (size_t)blockIdx.x * 96 * kGramLd;
    switch (warp) {
    case 0: gram_store<0>(acc, P, lane); break;
    case 1: gram_store<1>(acc, P, lane); break;
    case 2: gram_store<2>(acc, P, lane); break;
    case 3: gram_store<3>(acc, P, lane); break;
    case 4: gram_store<4>(acc, P, lane); break;
    default: gram_store<5>(acc, P, lane); break;
    }
}

// G[i][j] (k x k, both triangles) and b[i] from the partials: blockDim = (32 outputs, 8 slices of
// the partials); every slice is summed in order, then the 8 slice sums in order (deterministic)
__global__ void __launch_bounds__(256) gram_reduce_kernel(const double *partial, int nparts, int k, double *G, double *b) {
    BFD_SHARED_ARRAY(double, part, [8][33]);
    const int idx = blockIdx.x * 32 + threadIdx.x;
    const bool ok = idx < k * (k + 1);
    const int i = ok ? idx / (k + 1) : 0, j = ok ? idx - i * (k + 1) : 0;
    int r = i, c = j;
    if (j == k) c = 96;                                      // the f column
    else if ((i >> 3) > (j >> 3)) { r = j; c = i; }           // below the block diagonal: mirror
    double s = 0.0;
    if (ok)
        for (int p = threadIdx.y; p < nparts; p += 8) s += partial[((size_t)p * 96 + r) * kGramLd + c];
    part[threadIdx.y][threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.y == 0 && ok) {
        double t = 0.0;
#pragma unroll
        for (int q = 0; q < 8; ++q) t += part[q][threadIdx.x];
        if (j == k) b[i] = t; else G[(size_t)i * k + j] = t;
    }
}

}  // namespace bfd

using namespace bfd;

static int gram_grid(long long M) {
    const long long ntiles = (M + kGramRows - 1) / kGramRows;
    long long g = 148 * 3;                                    // 3 CTAs of 6 warps per SM are resident (94 registers)
    if (g > ntiles) g = ntiles;
    return (int)BFD_GRID((int)g);
}

extern "C" size_t bfd_gram_workspace_bytes(int64_t M, int k) {
    if (M < 1 || k < 1 || k > 96) return 0;
    return (size_t)gram_grid(M) * 96 * kGramLd * sizeof(double);
}

extern "C" int bfd_gram(const double *X, const double *f, int64_t M, int k, double *G, double *b, void *workspace,
                        size_t workspace_bytes, void *stream) {
    if (!X || !f || !G || !b || !workspace || M < 1 || k < 1 || k > 96) return BFD_EINVAL;
    if (workspace_bytes < bfd_gram_workspace_bytes(M, k)) return BFD_EWORKSPACE;
    GramArgs a{};
    a.X = X; a.f = f; a.M = M; a.k = k;
    a.ntiles = (int)((M + kGramRows - 1) / kGramRows);
    a.partial = (double *)workspace;
    const int grid = gram_grid(M);
    cudaStream_t st = (cudaStream_t)stream;
    BFD_LAUNCH(gram_partial_kernel, grid, 32 * kGramWarps, 0, st, a);
    BFD_LAUNCHED();
    BFD_LAUNCH(gram_reduce_kernel, (k * (k + 1) + 31) / 32, dim3(32, 8), 0, st, (const double *)a.partial, grid, k, G, b);
    BFD_LAUNCHED();
    return BFD_OK;
}

// ------------------------------------------------------------------------------------------
// bfd_gram_static: the same normal equations for callers that cannot allocate (CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The partial sums live in a file-scope device array of
// fixed size instead of a caller-sized workspace: BFD_GRAM_STATIC_PARTS partials (one CTA per SM)
// of 96 x 104 doubles = 11.3 MiB, held by every GPU that loads the library until the process
// exits.  In HOST code the array's name is only a stand-in for the symbol, not a device address:
// the address is looked up once per device with cudaGetSymbolAddress (that device current) and
// passed to the kernels like any other pointer.  Nothing has to start at zero: every CTA
// overwrites each entry of its partial that gram_reduce_kernel reads.  All calls share the one
// array, so calls that overlap on the GPU (different streams, no ordering between them) race.
// ------------------------------------------------------------------------------------------
namespace bfd {
static __device__ __align__(256) double g_gram_static_scratch[(size_t)BFD_GRAM_STATIC_PARTS * 96 * kGramLd];
constexpr int kMaxDevices = 64;
static double *g_gram_static_ptr[kMaxDevices];       // device address of the array on each device, 0 = not looked up yet

static int gram_static_scratch(double **out) {
    int dev = 0;
    BFD_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) {
        snprintf(g_cuda_error, sizeof(g_cuda_error), "bfd_gram_static: device ordinal %d not in [0, %d)", dev, kMaxDevices);
        return BFD_ETOOBIG;
    }
    double *p = __atomic_load_n(&g_gram_static_ptr[dev], __ATOMIC_ACQUIRE);
    if (!p) {                                        // threads that race here look up the same address
        BFD_CUDA(cudaGetSymbolAddress((void **)&p, g_gram_static_scratch));
        if (!p) return BFD_ECUDA;
        __atomic_store_n(&g_gram_static_ptr[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return BFD_OK;
}
}  // namespace bfd

extern "C" int bfd_gram_static(const double *X, const double *f, int64_t M, int k, double *G, double *b, void *stream) {
    if (!X || !f || !G || !b || M < 1 || k < 1) return BFD_EINVAL;
    if (M > (int64_t)BFD_GRAM_STATIC_MAX_ROWS || k > 96) {
        snprintf(g_cuda_error, sizeof(g_cuda_error),
                 "bfd_gram_static: M = %lld, k = %d beyond the fixed scratch's maximum (M <= %lld, k <= 96); "
                 "use bfd_gram with a caller-sized workspace", (long long)M, k, (long long)BFD_GRAM_STATIC_MAX_ROWS);
        return BFD_ETOOBIG;
    }
    double *scratch = nullptr;
    const int rc = gram_static_scratch(&scratch);
    if (rc != BFD_OK) return rc;
    GramArgs a{};
    a.X = X; a.f = f; a.M = M; a.k = k;
    a.ntiles = (int)((M + kGramRows - 1) / kGramRows);
    a.partial = scratch;
    int grid = a.ntiles < BFD_GRAM_STATIC_PARTS ? a.ntiles : BFD_GRAM_STATIC_PARTS;
    grid = (int)BFD_GRID(grid);
    cudaStream_t st = (cudaStream_t)stream;
    BFD_LAUNCH(gram_partial_kernel, grid, 32 * kGramWarps, 0, st, a);
    BFD_LAUNCHED();
    BFD_LAUNCH(gram_reduce_kernel, (k * (k + 1) + 31) / 32, dim3(32, 8), 0, st, (const double *)scratch, grid, k, G, b);
    BFD_LAUNCHED();
    return BFD_OK;
}

// ------------------------------------------------------------------------------------------
// G dy = b for the k x k normal equations (k <= 96): Cholesky factorisation and the triangular
// solves by ONE CTA, thread i owning row i of the packed factor in shared memory.  Replaces the
// LAPACK solve inside numpy.linalg.lstsq (:1994, :2038).  A library potrf + potrs on a 95 x 95
// matrix is a chain of about ten tiny kernels (~0.1 ms); this is one launch:
//   * left-looking Cholesky, ONE barrier per column: at column j every thread i >= j forms
//     a_ij - sum_p L_ip L_jp from its own row and row j (a broadcast read), and also the diagonal
//     a_jj - sum_p L_jp^2 itself rather than waiting for thread j to publish it;
//   * b rides along as row k of the matrix: the same recurrence then yields z = L^-1 b for free;
//   * back substitution L^T x = z column by column, one barrier per column.
// ------------------------------------------------------------------------------------------
namespace bfd {

constexpr int kSolveThreads = 128;

__device__ __forceinline__ int tri(int i, int j) { return i * (i + 1) / 2 + j; }      // j <= i

constexpr int kSolveNB = 8;           // columns per panel

// Blocked left-looking Cholesky (panels of 8 columns) + blocked back substitution, thread t = row t of the
// packed factor, row k = scale * b (so the factorisation also yields z = L^-1 b).  Per panel: every row
// subtracts the contribution of the finished columns from its 8 panel entries (8 dot products that share
// the row's loads), every thread factors the 8 x 8 diagonal block for itself in registers (no broadcast, no
// extra barrier) and solves its own row against it.  3 barriers per panel instead of one per column and
// 8-way instruction-level parallelism in the dot products: 71 us -> ~15 us at k = 95 (the column-by-column
// version spent its time in ~190 barriers and as many dependent rsqrt / FMA chains).
__global__ void __launch_bounds__(kSolveThreads) spd_solve_kernel(const double *G, const double *b, int k, double scale,
                                                                    double *x) {
    BFD_SHARED_ARRAY(double, L, [97 * 98 / 2]);        // rows 0..k-1: the factor; row k: scale*b, then z
    BFD_SHARED_ARRAY(double, xs, [104]);
    BFD_SHARED_ARRAY(double, dg, [104]);              // reciprocal diagonal of the factor
    const int t = threadIdx.x;
    const int kin = k;
    k = (kin + kSolveNB - 1) & ~(kSolveNB - 1);          // padded with an identity block: every panel is full
    for (int i = kin + (t >> 3); i < k; i += kSolveThreads >> 3)
        for (int j = t & 7; j <= i; j += 8) L[tri(i, j)] = i == j ? 1.0 : 0.0;
#pragma unroll 8
    for (int e = t; e < kin * kin; e += kSolveThreads) {
        const double g = G[e];                           // unconditional: the loads of 8 iterations go out together
        const int i = e / kin, j = e - i * kin;
        if (j <= i) L[tri(i, j)] = g;
    }
    if (t < k) L[tri(k, t)] = t < kin ? scale * b[t] : 0.0;
    __syncthreads();
    const bool row = t <= k;
    const int ri = tri(row ? t : k, 0);
    for (int j0 = 0; j0 < k; j0 += kSolveNB) {
        constexpr int nb = kSolveNB;
        double acc[kSolveNB];
        // phase 1: panel entries of this row minus the finished columns' contribution
        if (row && t >= j0) {
#pragma unroll
            for (int c = 0; c < kSolveNB; ++c) acc[c] = (c < nb && j0 + c <= t) ? L[ri + j0 + c] : 0.0;
            for (int q = 0; q < j0; ++q) {
                const double lt = L[ri + q];
#pragma unroll
                for (int c = 0; c < kSolveNB; ++c) {
                    const int rc = j0 + c < k ? j0 + c : k - 1;        // padded columns: any finished row, weight 0 below
                    acc[c] = fma(-lt, L[tri(rc, q)], acc[c]);
                }
            }
#pragma unroll
            for (int c = 0; c < kSolveNB; ++c)
                if (c < nb && j0 + c <= t) L[ri + j0 + c] = acc[c];
        }
        __syncthreads();
        // phase 2: the 8 x 8 diagonal block, factored by every thread for itself (registers)
        double D[kSolveNB][kSolveNB], rd[kSolveNB];
        if (row && t >= j0) {
#pragma unroll
            for (int c = 0; c < kSolveNB; ++c)
#pragma unroll
                for (int d = 0; d <= c; ++d) D[c][d] = (c < nb) ? L[tri(j0 + c, j0 + d)] : (c == d ? 1.0 : 0.0);
#pragma unroll
            for (int c = 0; c < kSolveNB; ++c) {
#pragma unroll
                for (int d = 0; d < c; ++d) {
                    double v = D[c][d];
#pragma unroll
                    for (int e = 0; e < d; ++e) v = fma(-D[c][e], D[d][e], v);
                    D[c][d] = v * rd[d];
                }
                double dd = D[c][c];
#pragma unroll
                for (int e = 0; e < c; ++e) dd = fma(-D[c][e], D[c][e], dd);
                rd[c] = rsqrt(dd);                              // NaN for a matrix that is not positive definite
            }
        }
        __syncthreads();
        // phase 3: rows below the block solve against it; rows of the block take their factored entries
        if (row && t >= j0) {
            if (t >= j0 + nb) {
#pragma unroll
                for (int c = 0; c < kSolveNB; ++c) {
                    double v = acc[c];
#pragma unroll
                    for (int d = 0; d < c; ++d) v = fma(-acc[d], D[c][d], v);
                    acc[c] = v * rd[c];
                    if (c < nb) L[ri + j0 + c] = acc[c];
                }
            } else {
                const int c = t - j0;
#pragma unroll
                for (int cc = 0; cc < kSolveNB; ++cc)
                    if (cc == c) {
#pragma unroll
                        for (int d = 0; d < cc; ++d) L[ri + j0 + d] = D[cc][d];
                        dg[t] = rd[cc];
                    }
            }
        }
        __syncthreads();
    }
    if (t < k) xs[t] = L[tri(k, t)];                              // z
    __syncthreads();
    // back substitution L^T x = z, a panel at a time: every thread solves the 8 x 8 triangle for itself,
    // rows above the panel subtract its contribution, rows of the panel keep their component
    for (int j0 = ((k - 1) / kSolveNB) * kSolveNB; j0 >= 0; j0 -= kSolveNB) {
        constexpr int nb = kSolveNB;
        double xb[kSolveNB];
        if (t < j0 + nb) {
#pragma unroll
            for (int c = kSolveNB - 1; c >= 0; --c) {
                double v = 0.0;
                if (c < nb) {
                    v = xs[j0 + c];
#pragma unroll
                    for (int d = kSolveNB - 1; d > c; --d)
                        if (d < nb) v = fma(-L[tri(j0 + d, j0 + c)], xb[d], v);
                    v *= dg[j0 + c];
                }
                xb[c] = v;
            }
        }
        __syncthreads();                                          // every thread has read the panel's z
        if (t < j0) {
            double v = xs[t];
#pragma unroll
            for (int c = 0; c < kSolveNB; ++c)
                if (c < nb) v = fma(-L[tri(j0 + c, t)], xb[c], v);
            xs[t] = v;
        } else if (t < j0 + nb) {
#pragma unroll
            for (int c = 0; c < kSolveNB; ++c)
                if (c == t - j0) xs[t] = xb[c];
        }
        __syncthreads();
    }
    if (t < kin) x[t] = xs[t];
}

}  // namespace bfd

extern "C" int bfd_spd_solve(const double *G, const double *b, int k, double scale, double *x, void *stream) {
    if (!G || !b || !x || k < 1 || k > 96) return BFD_EINVAL;
    BFD_LAUNCH(spd_solve_kernel, 1, kSolveThreads, 0, (cudaStream_t)stream, G, b, k, scale, x);
    BFD_LAUNCHED();
    return BFD_OK;
}

// ------------------------------------------------------------------------------------------
// w = V y (basis.dot(y), :1999 / :2042 / :196): M rows of k <= 96 doubles against a k-vector.
// HBM-bound: 8 M k bytes.  A warp takes 4 consecutive rows (3 KB contiguous, lane = mode, three
// chunks of 32 modes), keeps y in registers and reduces the four partial sums together: two
// "give half, keep half" exchanges leave every lane with one row's partial, three more finish
// it -- 6 shuffles for 4 rows instead of 20.
// ------------------------------------------------------------------------------------------
namespace bfd {

template <int MC>
__global__ void __launch_bounds__(256) basis_times_y_kernel(const double *__restrict__ V, const double *__restrict__ y,
                                                           long long M, int k, double *__restrict__ w) {
    const int lane = threadIdx.x & 31;
    double yr[MC];
    bool ok[MC];
#pragma unroll
    for (int m = 0; m < MC; ++m) {
        ok[m] = lane + 32 * m < k;
        yr[m] = ok[m] ? y[lane + 32 * m] : 0.0;
    }
    const long long nquads = (M + 3) / 4;
    const long long warp0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long q = warp0; q < nquads; q += nwarps) {
        const long long r0 = q * 4;
        // all 4 x MC loads go out before the first product: with the loads of a row issued next to
        // their DFMAs the compiler reuses three registers and only 768 B per warp are in flight
        // (3.6 TB/s instead of the 6.8 TB/s a read-only stream reaches on this part)
        double v[4][MC];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const long long r = r0 + i < M ? r0 + i : M - 1;      // tail rows: recomputed, not stored
            const double *row = V + r * k + lane;
#pragma unroll
            for (int m = 0; m < MC; ++m) v[i][m] = ok[m] ? __ldg(row + 32 * m) : 0.0;
        }
        double a[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            double s = 0.0;
#pragma unroll
            for (int m = 0; m < MC; ++m) s += v[i][m] * yr[m];
            a[i] = s;
        }
        const bool hi = lane & 16;
        double r0s = (hi ? a[2] : a[0]) + __shfl_xor_sync(0xffffffffu, hi ? a[0] : a[2], 16);
        double r1s = (hi ? a[3] : a[1]) + __shfl_xor_sync(0xffffffffu, hi ? a[1] : a[3], 16);
        const bool mid = lane & 8;
        double s = (mid ? r1s : r0s) + __shfl_xor_sync(0xffffffffu, mid ? r0s : r1s, 8);
        s += __shfl_xor_sync(0xffffffffu, s, 4);
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        const long long row = r0 + (hi ? 2 : 0) + (mid ? 1 : 0);
        if ((lane & 7) == 0 && row < M) w[row] = s;
    }
}

// The main kernel: the 4 rows of a quad are 4k contiguous doubles starting on a 16-byte boundary.
// Every warp streams its quads through a private 3-stage shared-memory ring with 16-byte cp.async
// copies, so the bytes in flight do not depend on how the compiler schedules register loads (the
// register version above keeps 768 B per warp in flight: ptxas interleaves loads and DFMAs to save
// registers, and the kernel stops at 3.6 TB/s where a read-only stream reaches 6.8 TB/s).  Which
// row and which mode an element belongs to depends only on (lane, pair), not on the quad, so the
// matching y values and row numbers are looked up once, before the loop.
constexpr int kVyWarps = 4, kVyStages = 3, kVyStageDoubles = 4 * 96;

template <int NP>                       // 16-byte pairs per lane: ceil(2k / 32)
__global__ void __launch_bounds__(32 * kVyWarps) basis_times_y_async_kernel(const double *__restrict__ V,
                                                                           const double *__restrict__ y, long long nquads,
                                                                           int k, double *__restrict__ w) {
    BFD_SHARED_ARRAY(double, ring, [kVyWarps][kVyStages][kVyStageDoubles]);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double yv[NP][2];
    int rid[NP][2];
#pragma unroll
    for (int i = 0; i < NP; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int e = 2 * (lane + 32 * i) + j;
            const bool ok = e < 4 * k;
            const int row = ok ? e / k : 4;
            rid[i][j] = row;
            yv[i][j] = ok ? y[e - row * k] : 0.0;
        }
    const saddr mine = smem_addr(&ring[warp][0][0]);
    const long long warp0 = (long long)blockIdx.x * kVyWarps + warp, nwarps = (long long)gridDim.x * kVyWarps;
    auto issue = [&](long long q, int st) {
        if (q < nquads) {
            const double *base = V + q * 4 * k;
#pragma unroll
            for (int i = 0; i < NP; ++i) {
                const int p = lane + 32 * i;
                if (2 * p < 4 * k) cp_async16(mine + (saddr)((st * kVyStageDoubles + 2 * p) * 8), base + 2 * p);
            }
        }
        cp_async_commit();
    };
    for (int s = 0; s < kVyStages - 1; ++s) issue(warp0 + s * nwarps, s);
    int it = 0;
    for (long long q = warp0; q < nquads; q += nwarps, ++it) {
        cp_async_wait_pending<kVyStages - 2>();
        __syncwarp();                                   // every lane's copies of this quad have landed; last quad's reads are done
        issue(q + (kVyStages - 1) * nwarps, (it + kVyStages - 1) % kVyStages);
        const saddr sb = mine + (saddr)((it % kVyStages) * kVyStageDoubles * 8);
        double a[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int p = lane + 32 * i;
            double v0 = 0.0, v1 = 0.0;
            if (2 * p < 4 * k) lds2(sb + (saddr)(2 * p * 8), v0, v1);
            const double p0 = v0 * yv[i][0], p1 = v1 * yv[i][1];
#pragma unroll
            for (int r = 0; r < 4; ++r) a[r] += (rid[i][0] == r ? p0 : 0.0) + (rid[i][1] == r ? p1 : 0.0);
        }
        const bool hi = lane & 16;
        double r0s = (hi ? a[2] : a[0]) + __shfl_xor_sync(0xffffffffu, hi ? a[0] : a[2], 16);
        double r1s = (hi ? a[3] : a[1]) + __shfl_xor_sync(0xffffffffu, hi ? a[1] : a[3], 16);
        const bool mid = lane & 8;
        double s = (mid ? r1s : r0s) + __shfl_xor_sync(0xffffffffu, mid ? r0s : r1s, 8);
        s += __shfl_xor_sync(0xffffffffu, s, 4);
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        if ((lane & 7) == 0) w[q * 4 + (hi ? 2 : 0) + (mid ? 1 : 0)] = s;
    }
}

}  // namespace bfd

static int basis_times_y_rows(const double *V, const double *y, long long M, int k, double *w, cudaStream_t st) {
    long long blocks = ((M + 3) / 4 + 7) / 8;
    if (blocks > 148 * 8) blocks = 148 * 8;
    const int grid = (int)BFD_GRID((int)blocks);
    if (k > 64) BFD_LAUNCH((basis_times_y_kernel<3>), grid, 256, 0, st, V, y, M, k, w);
    else if (k > 32) BFD_LAUNCH((basis_times_y_kernel<2>), grid, 256, 0, st, V, y, M, k, w);
    else BFD_LAUNCH((basis_times_y_kernel<1>), grid, 256, 0, st, V, y, M, k, w);
    BFD_LAUNCHED();
    return BFD_OK;
}

extern "C" int bfd_basis_times_y(const double *V, const double *y, int64_t M, int k, double *w, void *stream) {
    if (!V || !y || !w || M < 1 || k < 1 || k > 96) return BFD_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    const long long nquads = ((uintptr_t)V & 15) ? 0 : M / 4;      // the ring kernel copies 16 bytes at a time
    if (nquads > 0) {
        long long blocks = (nquads + kVyWarps - 1) / kVyWarps;
        if (blocks > 148 * 5) blocks = 148 * 5;             // 89 registers, 36.9 KB of shared memory: 5 CTAs per SM
        const int grid = (int)BFD_GRID((int)blocks);
        switch ((2 * k + 31) / 32) {
        case 1: BFD_LAUNCH((basis_times_y_async_kernel<1>), grid, 32 * kVyWarps, 0, st, V, y, nquads, k, w); break;
        case 2: BFD_LAUNCH((basis_times_y_async_kernel<2>), grid, 32 * kVyWarps, 0, st, V, y, nquads, k, w); break;
        case 3: BFD_LAUNCH((basis_times_y_async_kernel<3>), grid, 32 * kVyWarps, 0, st, V, y, nquads, k, w); break;
        case 4: BFD_LAUNCH((basis_times_y_async_kernel<4>), grid, 32 * kVyWarps, 0, st, V, y, nquads, k, w); break;
        case 5: BFD_LAUNCH((basis_times_y_async_kernel<5>), grid, 32 * kVyWarps, 0, st, V, y, nquads, k, w); break;
        default: BFD_LAUNCH((basis_times_y_async_kernel<6>), grid, 32 * kVyWarps, 0, st, V, y, nquads, k, w); break;
        }
        BFD_LAUNCHED();
    }
    const long long done = nquads * 4;
    if (done < M) return basis_times_y_rows(V + done * k, y, M - done, k, w + done, st);   // tail rows (or everything)
    return BFD_OK;
}
