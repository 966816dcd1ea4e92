// dmma_rate.cu -- throughput and latency of the FP64 MMA (mma.sync m8n8k4 f64 = SASS DMMA.8x8x4) on this part, next to the
// plain DFMA rate: the numbers behind the choice of the DMMA reciprocal sweep (kernels_sweep.cuh, recip_groups).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench/dmma_rate tools/ubench/dmma_rate.cu && tools/ubench/dmma_rate
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
template <int CHAINS>
__global__ void k_dmma(double* out, int iters) {
    double c0[CHAINS], c1[CHAINS];
    const double a = 1.0 + threadIdx.x * 1e-9, b = 0.5;
#pragma unroll
    for (int i = 0; i < CHAINS; i++) { c0[i] = i; c1[i] = -i; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < CHAINS; i++) dmma(c0[i], c1[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; i++) s += c0[i] + c1[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int CHAINS>
__global__ void k_dfma(double* out, int iters) {
    double c[CHAINS];
    const double a = 0.999999 + threadIdx.x * 1e-12, b = 1e-7;
#pragma unroll
    for (int i = 0; i < CHAINS; i++) c[i] = i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < CHAINS; i++) c[i] = fma(c[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; i++) s += c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <typename F>
static float time_ms(F launch) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    return best;
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount, iters = 20000;
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    double* d; cudaMalloc(&d, (size_t)sms * 8 * 1024 * sizeof(double));
    printf("# %s, %d SMs, max clock %.0f MHz\n", p.name, sms, clk_khz / 1e3);
    for (int warps_per_sm : {4, 8, 16, 32}) {
        const int blocks = sms * (warps_per_sm / 4 > 0 ? warps_per_sm / 4 : 1), threads = 128;
        float ms = time_ms([&] { k_dmma<4><<<blocks, threads>>>(d, iters); });
        const double n = (double)blocks * (threads / 32) * iters * 4;
        printf("DMMA.8x8x4  %2d warps/SM, 4 independent chains: %.1f G DMMA/s = %.2f TFLOP/s (256 FMA each), %.2f cycles per DMMA per SM\n",
               warps_per_sm, n / ms / 1e6, n * 512 / ms / 1e9, ms * 1e-3 * clk_khz * 1e3 * sms / n);
    }
    {
        float ms = time_ms([&] { k_dmma<1><<<sms, 32>>>(d, iters); });
        printf("DMMA.8x8x4  dependent chain, 1 warp/SM: %.1f cycles latency\n", ms * 1e-3 * clk_khz * 1e3 / iters);
    }
    {
        const int blocks = sms * 8, threads = 256;
        float ms = time_ms([&] { k_dfma<8><<<blocks, threads>>>(d, iters); });
        const double n = (double)blocks * threads * iters * 8;
        printf("DFMA        64 warps/SM, 8 chains: %.2f TFLOP/s, %.2f cycles per warp-DFMA per SM\n", 2 * n / ms / 1e9, ms * 1e-3 * clk_khz * 1e3 * sms / (n / 32));
    }
    cudaFree(d);
    return 0;
}
