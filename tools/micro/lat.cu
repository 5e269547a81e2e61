// Microbenchmark: issue rate of DFMA / FFMA2 from ONE warp per SM sub-partition as a function of
// the number of independent dependency chains (gives latency and single-warp throughput), and
// with 2 / 3 / 4 warps per sub-partition.
#include <cstdio>
#include <cuda_runtime.h>

template <int C> __global__ void kd(double *out, long long *cyc, int iters, double s) {
    double f[C];
    for (int i = 0; i < C; ++i) f[i] = threadIdx.x * 1e-3 + i;
    double a = s, b = s * 0.5;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int i = 0; i < C; ++i) f[i] = fma(f[i], a, b);
    }
    long long t1 = clock64();
    double t = 0;
    for (int i = 0; i < C; ++i) t += f[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long r;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(*(unsigned long long *)&a), "l"(*(unsigned long long *)&b), "l"(*(unsigned long long *)&c));
    return *(float2 *)&r;
}
template <int C> __global__ void kf(float *out, long long *cyc, int iters, float s) {
    float2 f[C];
    for (int i = 0; i < C; ++i) f[i] = make_float2(threadIdx.x * 1e-3f + i, i);
    float2 a = make_float2(s, s), b = make_float2(s * 0.5f, s);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int i = 0; i < C; ++i) f[i] = ffma2(f[i], a, b);
    }
    long long t1 = clock64();
    float t = 0;
    for (int i = 0; i < C; ++i) t += f[i].x + f[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

int main() {
    double *o64; float *o32; long long *cyc, h;
    cudaMalloc(&o64, 148 * 1024 * 8); cudaMalloc(&o32, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
    const int iters = 20000;
#define RUN(K, C, W) { K<C><<<148, 128 * W>>>(K == kd<C> ? (decltype(o64))o64 : (decltype(o64))o64, cyc, iters, 1.0000001); }
#define RD(C, W) { kd<C><<<148, 128 * W>>>(o64, cyc, iters, 1.0000001); cudaDeviceSynchronize(); kd<C><<<148, 128 * W>>>(o64, cyc, iters, 1.0000001); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); \
    printf("DFMA  chains=%2d warps/SMSP=%d: %.2f cycles per DFMA per warp, %.2f cycles per DFMA per SMSP\n", C, W, (double)h / (iters * 4.0 * C), (double)h / (iters * 4.0 * C * W)); }
#define RF(C, W) { kf<C><<<148, 128 * W>>>(o32, cyc, iters, 1.0000001f); cudaDeviceSynchronize(); kf<C><<<148, 128 * W>>>(o32, cyc, iters, 1.0000001f); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); \
    printf("FFMA2 chains=%2d warps/SMSP=%d: %.2f cycles per FFMA2 per warp, %.2f cycles per FFMA2 per SMSP\n", C, W, (double)h / (iters * 4.0 * C), (double)h / (iters * 4.0 * C * W)); }
    RD(1, 1) RD(2, 1) RD(3, 1) RD(4, 1) RD(6, 1) RD(8, 1) RD(12, 1) RD(16, 1)
    RD(1, 2) RD(2, 2) RD(4, 2) RD(8, 2) RD(2, 3) RD(4, 3) RD(2, 4) RD(4, 4)
    RF(1, 1) RF(2, 1) RF(3, 1) RF(4, 1) RF(6, 1) RF(8, 1) RF(12, 1) RF(16, 1)
    RF(1, 2) RF(2, 2) RF(4, 2) RF(8, 2) RF(2, 3) RF(4, 3) RF(2, 4) RF(4, 4)
    return 0;
}
