// Which SM sub-partition does warp w of a 5-warp CTA run on, with 4 CTAs resident per SM?
// (profiling aid for the role -> warp assignment of fom_chunk_kernel: the data-movement warp and a
// level warp sharing one scheduler would make that scheduler the bottleneck of the whole SM).
// Records %smid, %warpid (hardware warp slot) and %nwarpid for every warp of every CTA of a launch
// shaped like the hot kernel (160 threads, 44 KB of static shared memory, <= 4 CTAs per SM), and
// measures the slot -> scheduler mapping directly: two warps spinning on dependent DFMA chains slow
// each other down only when they share a scheduler's FP64 pipe.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/warp_slot_probe tools/warp_slot_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <map>
#include <vector>

__global__ void __launch_bounds__(160, 4) probe(unsigned *out, int hold) {
    __shared__ double pad[44 * 1024 / 8];
    unsigned smid, warpid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    asm volatile("mov.u32 %0, %%warpid;" : "=r"(warpid));
    pad[threadIdx.x] = 1.0;
    __syncthreads();
    // stay resident long enough for the whole first wave to be co-resident
    long long t0 = clock64();
    while (clock64() - t0 < hold) { }
    if ((threadIdx.x & 31) == 0) {
        unsigned *o = out + (blockIdx.x * 5 + (threadIdx.x >> 5)) * 2;
        o[0] = smid; o[1] = warpid;
    }
    if (pad[threadIdx.x] < 0) out[0] = 0;
}

// warps `a` and `b` of one 32-warp CTA (one CTA per SM -> slots 0..31 in order) run an ILP-4 DFMA
// loop, all other warps exit: the pair takes longer when both sit on one scheduler
__global__ void __launch_bounds__(1024, 1) pair(double *sink, long long *cyc, int a, int b, int n) {
    const int w = threadIdx.x >> 5;
    unsigned warpid;
    asm volatile("mov.u32 %0, %%warpid;" : "=r"(warpid));
    if (w != a && w != b) return;
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    const long long t0 = clock64();
    for (int i = 0; i < n; ++i) {
        x0 = fma(x0, 1.0000001, 0.5); x1 = fma(x1, 1.0000001, 0.5); x2 = fma(x2, 1.0000001, 0.5); x3 = fma(x3, 1.0000001, 0.5);
        x4 = fma(x4, 1.0000001, 0.5); x5 = fma(x5, 1.0000001, 0.5); x6 = fma(x6, 1.0000001, 0.5); x7 = fma(x7, 1.0000001, 0.5);
    }
    const long long t1 = clock64();
    sink[threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
    if ((threadIdx.x & 31) == 0 && w == a) { cyc[0] = t1 - t0; cyc[1] = warpid; }
    if ((threadIdx.x & 31) == 0 && w == b) { cyc[2] = t1 - t0; cyc[3] = warpid; }
}

int main() {
    const int grid = 148 * 4;
    unsigned *d, *h = (unsigned *)malloc(grid * 5 * 2 * sizeof(unsigned));
    cudaMalloc(&d, grid * 5 * 2 * sizeof(unsigned));
    probe<<<grid, 160>>>(d, 2000000);
    cudaMemcpy(h, d, grid * 5 * 2 * sizeof(unsigned), cudaMemcpyDeviceToHost);
    printf("probe status: %s\n", cudaGetErrorString(cudaGetLastError()));
    // per SM: the warp slots of each resident CTA
    std::map<unsigned, std::vector<int>> ctas_of_sm;
    for (int c = 0; c < grid; ++c) ctas_of_sm[h[c * 10]].push_back(c);
    int shown = 0;
    long long hist[4][5] = {};   // [slot % 4][warp of CTA]
    for (auto &kv : ctas_of_sm) {
        if (shown < 4) {
            printf("SM %u:", kv.first);
            for (int c : kv.second) {
                printf("  CTA %d slots", c);
                for (int w = 0; w < 5; ++w) printf(" %u", h[(c * 5 + w) * 2 + 1]);
            }
            printf("\n");
            ++shown;
        }
        for (int c : kv.second)
            for (int w = 0; w < 5; ++w) ++hist[h[(c * 5 + w) * 2 + 1] % 4][w];
    }
    printf("CTAs per SM: min/max over SMs =");
    size_t mn = 99, mx = 0;
    for (auto &kv : ctas_of_sm) { mn = kv.second.size() < mn ? kv.second.size() : mn; mx = kv.second.size() > mx ? kv.second.size() : mx; }
    printf(" %zu / %zu on %zu SMs\n", mn, mx, ctas_of_sm.size());
    printf("count of (slot %% 4, CTA-local warp):\n");
    for (int s = 0; s < 4; ++s) { printf("  slot%%4=%d:", s); for (int w = 0; w < 5; ++w) printf(" w%d:%lld", w, hist[s][w]); printf("\n"); }

    // slot -> scheduler: time warp 0 together with warp b
    double *sink; long long *cyc, hc[4];
    cudaMalloc(&sink, 1024 * sizeof(double));
    cudaMalloc(&cyc, 4 * sizeof(long long));
    for (int b = 1; b < 12; ++b) {
        pair<<<1, 1024>>>(sink, cyc, 0, b, 200000);
        cudaMemcpy(hc, cyc, sizeof(hc), cudaMemcpyDeviceToHost);
        printf("warps 0 (slot %lld) + %d (slot %lld): %.3f / %.3f cycles per 8 DFMA\n", hc[1], b, hc[3], hc[0] / 200000.0, hc[2] / 200000.0);
    }
    printf("status: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
