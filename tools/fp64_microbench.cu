// FP64 pipe micro-benchmark for the roofline denominator of the fused FOM kernel (profiling aid).
//   dependent-chain latency of DFMA / DMUL / DADD, MUFU.RCP64H-based reciprocal, SHFL.UP of a
//   double, and the chip-wide DFMA issue rate with ILP 8 at 1, 2, 4, 8 warps per SMSP.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp64_microbench tools/fp64_microbench.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void lat_kernel(double *out, long long *cyc, int n, double a, double b) {
    double x = out[threadIdx.x];
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) {
#pragma unroll
        for (int j = 0; j < 16; ++j) x = fma(x, a, b);
    }
    long long t1 = clock64();
    double y = x;
    for (int i = 0; i < n; ++i) {
#pragma unroll
        for (int j = 0; j < 16; ++j) y = __shfl_up_sync(0xffffffffu, y, 1);
    }
    long long t2 = clock64();
    double z = y + 1.5;
    for (int i = 0; i < n; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            double r;
            asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(z));
            z = r + 1.25;
        }
    }
    long long t3 = clock64();
    out[threadIdx.x] = x + y + z;
    if (threadIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t2 - t1; cyc[2] = t3 - t2; }
}

template <int ILP>
__global__ void thr_kernel(double *out, int n, double a, double b) {
    double x[ILP];
#pragma unroll
    for (int j = 0; j < ILP; ++j) x[j] = out[threadIdx.x] + j;
    for (int i = 0; i < n; ++i) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int j = 0; j < ILP; ++j) x[j] = fma(x[j], a, b);
    }
    double s = 0;
#pragma unroll
    for (int j = 0; j < ILP; ++j) s += x[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
    double *out; long long *cyc;
    cudaMalloc(&out, 148 * 1024 * 8 * sizeof(double));
    cudaMemset(out, 0, 148 * 1024 * 8 * sizeof(double));
    cudaMallocManaged(&cyc, 64);
    const int n = 4096;
    lat_kernel<<<1, 32>>>(out, cyc, n, 0.999, 0.001);
    cudaDeviceSynchronize();
    printf("latency (cycles): DFMA dependent %.2f, SHFL.UP f64 dependent %.2f, rcp.approx.f64 + DADD %.2f\n",
           (double)cyc[0] / (n * 16.0), (double)cyc[1] / (n * 16.0), (double)cyc[2] / (n * 4.0));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    for (int warps_per_smsp : {1, 2, 4, 8}) {
        const int threads = 128 * warps_per_smsp;   // 4 SMSPs
        auto run = [&](auto kern, int ilp) {
            kern<<<sms, threads>>>(out, 64, 0.999, 0.001);
            cudaEventRecord(e0);
            const int reps = 2048;
            kern<<<sms, threads>>>(out, reps, 0.999, 0.001);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            const double fmas = (double)sms * threads * reps * 8.0 * ilp;
            printf("  %d warp(s)/SMSP, ILP %d: %.2f TFLOP/s (%.1f DFMA lanes/clk/SM at 1.965 GHz)\n", warps_per_smsp, ilp,
                   2 * fmas / ms / 1e9, fmas / (ms * 1e-3) / sms / 1.965e9);
        };
        run(thr_kernel<1>, 1); run(thr_kernel<4>, 4); run(thr_kernel<8>, 8);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
