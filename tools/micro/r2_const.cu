// Round-2 microbenchmark B: the FP64 FMA rate when one multiplicand is WARP-UNIFORM and comes from the
// constant bank (no shared-memory load, no register) -- the shape of a "one lane per row" W update, where
// every lane of a warp needs the same H[a][j] at the same time.  Streams S doubles of __constant__ memory
// per pass (S = 228: one 38-column chunk of H at k = 6; 1812: all of H at 302 columns; 8190: 64 KB).
// Compared with the same arithmetic on register operands.  Prints cycles per DFMA per SM sub-partition.
#include <cstdio>
#include <cuda_runtime.h>

__constant__ double Hc[8192];

// 12 independent accumulators (num[6], den[6]); per column j: 6 constants, each used twice
template <int MODE>
__global__ void __launch_bounds__(512) k(double *out, long long *cyc, int cols, int passes, double s) {
    double num[6], den[6];
#pragma unroll
    for (int a = 0; a < 6; ++a) { num[a] = a * 1e-3 + threadIdx.x; den[a] = -a * 1e-3; }
    double x = s + threadIdx.x * 1e-6, y = s * 0.5;
    double hr[6] = {s, s + 1, s + 2, s + 3, s + 4, s + 5};
    __syncthreads();
    const long long t0 = clock64();
    for (int p = 0; p < passes; ++p) {
#pragma unroll 4
        for (int j = 0; j < cols; ++j) {
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                const double h = MODE == 0 ? hr[a] : Hc[j * 6 + a];
                num[a] = fma(x, h, num[a]);
                den[a] = fma(y, h, den[a]);
            }
        }
    }
    const long long t1 = clock64();
    double t = 0;
#pragma unroll
    for (int a = 0; a < 6; ++a) t += num[a] + den[a];
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE> void run(const char *name, int cols, int nw, double *out, long long *cyc, int sms) {
    const int passes = 200000 / cols + 1;
    k<MODE><<<sms, nw * 32>>>(out, cyc, cols, passes, 1.0);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
    long long h[256];
    cudaMemcpy(h, cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost);
    double mean = 0;
    for (int i = 0; i < sms; ++i) mean += (double)h[i];
    mean /= sms;
    const double dfma_per_smsp = (double)passes * cols * 12 * nw / 4.0;
    printf("%-34s doubles streamed/pass=%5d warps/SM=%2d: %5.2f cycles per DFMA per sub-partition\n", name, cols * 6, nw,
           mean / dfma_per_smsp);
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    printf("%s, %d SMs\n", p.name, sms);
    double *out;
    long long *cyc;
    cudaMalloc(&out, sms * 512 * 8);
    cudaMalloc(&cyc, sms * 8);
    static double h[8192];
    for (int i = 0; i < 8192; ++i) h[i] = 1.0 + 1e-9 * i;
    cudaMemcpyToSymbol(Hc, h, sizeof(h));
    for (int nw : {4, 8, 12}) {
        run<0>("register operands", 38, nw, out, cyc, sms);
        for (int cols : {38, 302, 1365}) run<1>("constant-bank operand", cols, nw, out, cyc, sms);
    }
    return 0;
}
