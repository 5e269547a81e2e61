// Microbenchmark: do the legacy tensor-core paths (DMMA m8n8k4, TF32 m16n8k8) run
// concurrently with the FP64 / FP32 FMA pipes on B200?  Prints MAC/clk/SM for each mix.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void tf32mma(float (&c)[4], unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned b0, unsigned b1) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long r;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(*(unsigned long long *)&a), "l"(*(unsigned long long *)&b), "l"(*(unsigned long long *)&c));
    return *(float2 *)&r;
}

template <int NF, int NM>   // per loop iteration: NF x 8 DFMA and NM x 8 DMMA (independent chains)
__global__ void k64(double *out, int iters, double s) {
    double f[8], c0[8], c1[8];
    for (int i = 0; i < 8; ++i) { f[i] = threadIdx.x * 1e-3 + i; c0[i] = i; c1[i] = -i; }
    double a = s + threadIdx.x, b = s * 0.5;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < (NF > NM ? NF : NM); ++r) {
            if (r < NF) {
#pragma unroll
                for (int i = 0; i < 8; ++i) f[i] = fma(f[i], a, b);
            }
            if (r < NM) {
#pragma unroll
                for (int i = 0; i < 8; ++i) dmma(c0[i], c1[i], a, b);
            }
        }
    }
    double t = 0;
    for (int i = 0; i < 8; ++i) t += f[i] + c0[i] + c1[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
}

template <int NF, int NM>
__global__ void k32(float *out, int iters, float s) {
    float2 f[8];
    float c[8][4];
    for (int i = 0; i < 8; ++i) { f[i] = make_float2(threadIdx.x * 1e-3f + i, i); for (int j = 0; j < 4; ++j) c[i][j] = i + j; }
    float2 a = make_float2(s + threadIdx.x, s), b = make_float2(s * 0.5f, s);
    unsigned ua = __float_as_uint(s), ub = __float_as_uint(s * 0.25f);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < (NF > NM ? NF : NM); ++r) {
            if (r < NF) {
#pragma unroll
                for (int i = 0; i < 8; ++i) f[i] = ffma2(f[i], a, b);
            }
            if (r < NM) {
#pragma unroll
                for (int i = 0; i < 8; ++i) tf32mma(c[i], ua, ub, ua, ub, ub, ua);
            }
        }
    }
    float t = 0;
    for (int i = 0; i < 8; ++i) t += f[i].x + f[i].y + c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
}

template <typename F> float timeit(F launch) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(); cudaDeviceSynchronize();
    cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount; int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    double *o64; float *o32; cudaMalloc(&o64, sms * 4 * 512 * 8); cudaMalloc(&o32, sms * 4 * 512 * 4);
    const int iters = 20000, T = 512, B = sms * 2;
    printf("%s sms=%d nominal clk=%d kHz; rates below are per SM per ms (divide by actual clock)\n", p.name, sms, clk);
#define R64(NF, NM) { float ms = timeit([&] { k64<NF, NM><<<B, T>>>(o64, iters, 1.0000001); }); \
    double dfma = (double)B * T * iters * NF * 8, dm = (double)B * (T / 32) * iters * NM * 8 * 256; \
    printf("f64  DFMAx%d DMMAx%d: %.3f ms  DFMA %.1f MAC/ns/SM  DMMA %.1f MAC/ns/SM\n", NF, NM, ms, dfma / ms / 1e6 / sms, dm / ms / 1e6 / sms); }
    R64(1, 0) R64(0, 1) R64(1, 1) R64(2, 1) R64(4, 1) R64(8, 1)
#define R32(NF, NM) { float ms = timeit([&] { k32<NF, NM><<<B, T>>>(o32, iters, 1.0000001f); }); \
    double ff = (double)B * T * iters * NF * 8 * 2, tm = (double)B * (T / 32) * iters * NM * 8 * 1024; \
    printf("f32  FFMA2x%d TF32MMAx%d: %.3f ms  FFMA %.1f MAC/ns/SM  TF32 %.1f MAC/ns/SM\n", NF, NM, ms, ff / ms / 1e6 / sms, tm / ms / 1e6 / sms); }
    R32(1, 0) R32(0, 1) R32(1, 1) R32(2, 1) R32(4, 1) R32(8, 1)
    return 0;
}
