// Microbenchmark: does the DFMA / FFMA2 issue rate depend on how many distinct register operands an
// instruction reads (register-file bandwidth / reuse cache)?  One warp per sub-partition, 12 chains.
#include <cstdio>
#include <cuda_runtime.h>
#define N 12
// pattern 0: acc_i = fma(acc_i, a, b)      (1 new operand pair, 2 reused)
// pattern 1: acc_i = fma(x_i, c, acc_i)    (2 new, 1 reused)
// pattern 2: acc_i = fma(x_i, y_i, acc_i)  (3 new)
// pattern 3: acc_i = fma(x_i, y_(i+1), acc_i) with x, y rotating (3 new, no reuse possible)
template <int PAT> __global__ void kd(double *out, long long *cyc, int iters, double s) {
    double acc[N], x[N], y[N];
    for (int i = 0; i < N; ++i) { acc[i] = threadIdx.x * 1e-3 + i; x[i] = s + i * 1e-9; y[i] = s - i * 1e-9; }
    double a = s, b = s * 0.5;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int i = 0; i < N; ++i) {
                if (PAT == 0) acc[i] = fma(acc[i], a, b);
                if (PAT == 1) acc[i] = fma(x[i], a, acc[i]);
                if (PAT == 2) acc[i] = fma(x[i], y[i], acc[i]);
                if (PAT == 3) acc[i] = fma(x[(i + r) % N], y[(i + 2 * r + 1) % N], acc[i]);
            }
    }
    long long t1 = clock64();
    double t = 0;
    for (int i = 0; i < N; ++i) t += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long r;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(*(unsigned long long *)&a), "l"(*(unsigned long long *)&b), "l"(*(unsigned long long *)&c));
    return *(float2 *)&r;
}
template <int PAT> __global__ void kf(float *out, long long *cyc, int iters, float s) {
    float2 acc[N], x[N], y[N];
    for (int i = 0; i < N; ++i) { acc[i] = make_float2(threadIdx.x * 1e-3f + i, i); x[i] = make_float2(s + i * 1e-4f, s); y[i] = make_float2(s - i * 1e-4f, s); }
    float2 a = make_float2(s, s), b = make_float2(s * 0.5f, s);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int i = 0; i < N; ++i) {
                if (PAT == 0) acc[i] = ffma2(acc[i], a, b);
                if (PAT == 1) acc[i] = ffma2(x[i], a, acc[i]);
                if (PAT == 2) acc[i] = ffma2(x[i], y[i], acc[i]);
                if (PAT == 3) acc[i] = ffma2(x[(i + r) % N], y[(i + 2 * r + 1) % N], acc[i]);
            }
    }
    long long t1 = clock64();
    float t = 0;
    for (int i = 0; i < N; ++i) t += acc[i].x + acc[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
// scalar FFMA, two per "pack": is a pair of FFMAs cheaper than one FFMA2 when all three operands are new?
template <int PAT> __global__ void ks(float *out, long long *cyc, int iters, float s) {
    float2 acc[N], x[N], y[N];
    for (int i = 0; i < N; ++i) { acc[i] = make_float2(threadIdx.x * 1e-3f + i, i); x[i] = make_float2(s + i * 1e-4f, s); y[i] = make_float2(s - i * 1e-4f, s); }
    float2 a = make_float2(s, s), b = make_float2(s * 0.5f, s);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int i = 0; i < N; ++i) {
                if (PAT == 0) { acc[i].x = fmaf(acc[i].x, a.x, b.x); acc[i].y = fmaf(acc[i].y, a.y, b.y); }
                if (PAT == 1) { acc[i].x = fmaf(x[i].x, a.x, acc[i].x); acc[i].y = fmaf(x[i].y, a.y, acc[i].y); }
                if (PAT == 2) { acc[i].x = fmaf(x[i].x, y[i].x, acc[i].x); acc[i].y = fmaf(x[i].y, y[i].y, acc[i].y); }
                if (PAT == 3) {
                    acc[i].x = fmaf(x[(i + r) % N].x, y[(i + 2 * r + 1) % N].x, acc[i].x);
                    acc[i].y = fmaf(x[(i + r) % N].y, y[(i + 2 * r + 1) % N].y, acc[i].y);
                }
            }
    }
    long long t1 = clock64();
    float t = 0;
    for (int i = 0; i < N; ++i) t += acc[i].x + acc[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
    double *o64; float *o32; long long *cyc, h;
    cudaMalloc(&o64, 148 * 1024 * 8); cudaMalloc(&o32, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
    const int iters = 20000;
#define RD(P, W) { kd<P><<<148, 128 * W>>>(o64, cyc, iters, 1.0000001); cudaDeviceSynchronize(); kd<P><<<148, 128 * W>>>(o64, cyc, iters, 1.0000001); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); \
    printf("DFMA  pattern %d warps/SMSP=%d: %.2f cycles per DFMA per SMSP\n", P, W, (double)h / (iters * 4.0 * N * W)); }
#define RF(P, W) { kf<P><<<148, 128 * W>>>(o32, cyc, iters, 1.0000001f); cudaDeviceSynchronize(); kf<P><<<148, 128 * W>>>(o32, cyc, iters, 1.0000001f); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); \
    printf("FFMA2 pattern %d warps/SMSP=%d: %.2f cycles per FFMA2 per SMSP\n", P, W, (double)h / (iters * 4.0 * N * W)); }
#define RS(P, W) { ks<P><<<148, 128 * W>>>(o32, cyc, iters, 1.0000001f); cudaDeviceSynchronize(); ks<P><<<148, 128 * W>>>(o32, cyc, iters, 1.0000001f); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost); \
    printf("2xFFMA pattern %d warps/SMSP=%d: %.2f cycles per FFMA pair per SMSP\n", P, W, (double)h / (iters * 4.0 * N * W)); }
    RS(0, 1) RS(1, 1) RS(2, 1) RS(3, 1) RS(0, 3) RS(1, 3) RS(2, 3) RS(3, 3)
    RD(0, 1) RD(1, 1) RD(2, 1) RD(3, 1) RD(0, 3) RD(1, 3) RD(2, 3) RD(3, 3)
    RF(0, 1) RF(1, 1) RF(2, 1) RF(3, 1) RF(0, 3) RF(1, 3) RF(2, 3) RF(3, 3)
    return 0;
}
