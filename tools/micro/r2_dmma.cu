// Microbenchmark (round 2): mma.sync.m8n8k4.f64 (DMMA) from ONE warp per SM sub-partition --
//  (a) C independent accumulator chains (each DMMA depends on the previous one of its chain through C/D):
//      cycles per DMMA vs C gives the accumulate-to-accumulate latency and the single-warp issue rate;
//  (b) a chain through the A operand (D of one DMMA scaled by a DMUL is A of the next): the W H -> V*WH -> (V*WH) H^T
//      dependency of mma_kernel.cuh;
//  (c) 2 and 3 warps per sub-partition.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double &d0, double &d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
template <int C> __global__ void kacc(double *out, long long *cyc, int iters, double s) {
    double d0[C], d1[C];
    for (int i = 0; i < C; ++i) d0[i] = d1[i] = threadIdx.x * 1e-3 + i;
    const double a = s, b = s * 0.5;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int i = 0; i < C; ++i) dmma(d0[i], d1[i], a, b);
    }
    long long t1 = clock64();
    double t = 0;
    for (int i = 0; i < C; ++i) t += d0[i] + d1[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
// chain through A: d = mma(a, b, 0); a' = d0 * s  (DMUL)  ->  next mma
template <int C> __global__ void kchain(double *out, long long *cyc, int iters, double s) {
    double a[C];
    for (int i = 0; i < C; ++i) a[i] = threadIdx.x * 1e-3 + i;
    const double b = s * 0.5;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int i = 0; i < C; ++i) {
                double d0 = 0.0, d1 = 0.0;
                dmma(d0, d1, a[i], b);
                a[i] = d0 * s + d1;
            }
    }
    long long t1 = clock64();
    double t = 0;
    for (int i = 0; i < C; ++i) t += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

int main() {
    double *o; long long *cyc, h;
    cudaMalloc(&o, 148 * 1024 * 8); cudaMalloc(&cyc, 8);
    const int iters = 5000;
#define RA(C, W) { kacc<C><<<148, 128 * W>>>(o, cyc, iters, 1.0000001); cudaDeviceSynchronize(); kacc<C><<<148, 128 * W>>>(o, cyc, iters, 1.0000001); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); \
    printf("DMMA accumulate chains=%2d warps/SMSP=%d: %.2f cycles per DMMA per warp, %.2f per SMSP\n", C, W, (double)h / (iters * 4.0 * C), (double)h / (iters * 4.0 * C * W)); }
#define RC(C, W) { kchain<C><<<148, 128 * W>>>(o, cyc, iters, 1e-3); cudaDeviceSynchronize(); kchain<C><<<148, 128 * W>>>(o, cyc, iters, 1e-3); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); \
    printf("DMMA -> DFMA -> DMMA (through A) chains=%2d warps/SMSP=%d: %.2f cycles per step per warp, %.2f per SMSP\n", C, W, (double)h / (iters * 4.0 * C), (double)h / (iters * 4.0 * C * W)); }
    RA(1, 1) RA(2, 1) RA(3, 1) RA(4, 1) RA(6, 1) RA(8, 1)
    RA(1, 2) RA(2, 2) RA(4, 2) RA(1, 3) RA(2, 3) RA(4, 3)
    RC(1, 1) RC(2, 1) RC(3, 1) RC(4, 1) RC(6, 1) RC(1, 2) RC(2, 2) RC(1, 3) RC(2, 3)
    cudaError_t e = cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(e));
    return 0;
}
