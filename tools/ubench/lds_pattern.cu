// Shared-memory load cost by address pattern on sm_100a (cycles per warp-level LDS, all 4 SMSPs loaded).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o lds_pattern lds_pattern.cu && ./lds_pattern
#include <cstdio>
#include <cuda_runtime.h>
template <int W>  // W = 16: LDS.128 (double2), 8: LDS.64 (double)
__global__ void k(const int* idx, double* out, long long* cyc, int iters) {
    extern __shared__ __align__(16) unsigned char sm[];
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) ((double*)sm)[i] = i;
    __syncthreads();
    const int a = idx[threadIdx.x & 31];
    unsigned acc = 0;
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const int o = ((it * 8 + u) * 5) & 511;   // same shift for every lane: the pattern is preserved
            if (W == 16) { uint4 v = *(const uint4*)(sm + (size_t)(a + o) * 16); acc ^= v.x ^ v.y ^ v.z ^ v.w; }
            else { uint2 v = *(const uint2*)(sm + (size_t)(a + o) * 8); acc ^= v.x ^ v.y; }
        }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main() {
    const char* names[] = {"distinct consecutive", "uniform (1 address)", "2 addresses (half warps)", "3 addresses (11/11/10)", "2 addresses interleaved", "4 addresses (8 each)", "stride 2 elements", "MMA A fragment, row stride 36", "MMA A fragment, row stride 40"};
    int pat[9][32];
    for (int l = 0; l < 32; l++) { pat[0][l] = l; pat[1][l] = 5; pat[2][l] = l < 16 ? 3 : 40; pat[3][l] = l < 11 ? 3 : l < 22 ? 17 : 60; pat[4][l] = (l & 1) ? 3 : 40; pat[5][l] = 7 + 13 * (l / 8); pat[6][l] = 2 * l; pat[7][l] = (l & 3) * 36 + (l >> 2); pat[8][l] = (l & 3) * 40 + (l >> 2); }
    int* d_idx; double* d_out; long long* d_cyc;
    cudaMalloc(&d_idx, 128); cudaMalloc(&d_out, 1024 * 8 * 148); cudaMalloc(&d_cyc, 8 * 148);
    const int iters = 2000, threads = 512;
    for (int w = 0; w < 2; w++) for (int p = 0; p < 9; p++) {
        cudaMemcpy(d_idx, pat[p], 128, cudaMemcpyHostToDevice);
        if (w == 0) k<16><<<148, threads, 32768>>>(d_idx, d_out, d_cyc, iters); else k<8><<<148, threads, 32768>>>(d_idx, d_out, d_cyc, iters);
        long long c; cudaDeviceSynchronize(); cudaMemcpy(&c, d_cyc, 8, cudaMemcpyDeviceToHost);
        printf("%s  %-28s  %.2f cycles per warp-load (SM-wide: %d warps)\n", w == 0 ? "LDS.128" : "LDS.64 ", names[p], (double)c / (iters * 8.0 * (threads / 32)), threads / 32);
    }
    return 0;
}
