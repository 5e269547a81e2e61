// Round-2 microbenchmark A: what a shared-memory load costs the SM's LSU pipe as a function of the
// address pattern across the warp.  The question behind it: role W of the stream
// kernel loads the same H vector into several lanes -- does a broadcast save wavefronts, i.e. is the
// cost "bytes fetched from the banks" or "registers written"?
//
// Every warp issues ITER x 8 independent loads; all SMs run NW warps.  Prints SM cycles per warp-load.
#include <cstdio>
#include <cuda_runtime.h>

template <int BYTES> struct LdT;
template <> struct LdT<16> { typedef uint4 T; };
template <> struct LdT<8> { typedef uint2 T; };
template <> struct LdT<4> { typedef unsigned T; };

__device__ __forceinline__ unsigned fold(uint4 v) { return v.x ^ v.y ^ v.z ^ v.w; }
__device__ __forceinline__ unsigned fold(uint2 v) { return v.x ^ v.y; }
__device__ __forceinline__ unsigned fold(unsigned v) { return v; }

// off(lane) in bytes selects the pattern; `span` bytes of shared memory are walked with stride `step`
template <int BYTES>
__global__ void __launch_bounds__(512) lds_kernel(unsigned *out, long long *cyc, int iters, int mode, int span, int step) {
    extern __shared__ __align__(16) unsigned char sm[];
    typedef typename LdT<BYTES>::T T;
    for (int i = threadIdx.x; i < span / 4 + 8 * 1024 / 4; i += blockDim.x) ((unsigned *)sm)[i] = i * 2654435761u;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    int off;
    switch (mode) {
    case 0: off = lane * BYTES; break;                 // all lanes distinct, contiguous
    case 1: off = 0; break;                            // one address for the warp
    case 2: off = (lane & 7) * BYTES; break;           // 8 addresses, repeated in every quarter-warp
    case 3: off = (lane >> 3) * BYTES; break;          // 4 addresses, one per quarter-warp
    case 4: off = (lane >> 4) * BYTES; break;          // 2 addresses, one per half-warp
    case 5: off = lane * 304; break;                   // rows at a 304-byte pitch (a 38-column fp64 box)
    case 6: off = lane * 2416; break;                  // rows at the 302-column fp64 pitch
    case 7: off = (lane & 15) * BYTES; break;          // 16 addresses, repeated in both half-warps
    default: off = lane * BYTES; break;
    }
    const unsigned char *p = sm + off;
    unsigned acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int w = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) acc[u] ^= fold(*reinterpret_cast<const T *>(p + w + u * 1024));
        w += step;
        if (w >= span) w = 0;
    }
    const long long t1 = clock64();
    __syncthreads();
    unsigned a = 0;
    for (int u = 0; u < 8; ++u) a ^= acc[u];
    out[blockIdx.x * blockDim.x + threadIdx.x] = a;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int BYTES> void run(const char *name, int mode, int nw, unsigned *out, long long *cyc, int sms, int big) {
    const int iters = 4000;
    const int span = 4096, step = 16 * (BYTES / 4 > 0 ? 1 : 1);
    const int smem = (big ? 32 * 2416 + 16 : 0) + span + 8 * 1024 + 64;
    cudaFuncSetAttribute(lds_kernel<BYTES>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    lds_kernel<BYTES><<<sms, nw * 32, smem>>>(out, cyc, iters, mode, span, step);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
    long long h[256];
    cudaMemcpy(h, cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost);
    double mean = 0;
    for (int i = 0; i < sms; ++i) mean += (double)h[i];
    mean /= sms;
    printf("LDS.%-3d %-44s warps/SM=%2d: %6.2f SM cycles per warp-load\n", BYTES * 8, name, nw, mean / ((double)iters * 8 * nw));
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    printf("%s, %d SMs\n", p.name, sms);
    unsigned *out;
    long long *cyc;
    cudaMalloc(&out, sms * 512 * 4);
    cudaMalloc(&cyc, sms * 8);
    for (int nw : {4, 16}) {
        run<16>("32 distinct addresses (512 B)", 0, nw, out, cyc, sms, 0);
        run<16>("one address for the warp", 1, nw, out, cyc, sms, 0);
        run<16>("8 addresses, the same in each quarter-warp", 2, nw, out, cyc, sms, 0);
        run<16>("16 addresses, the same in each half-warp", 7, nw, out, cyc, sms, 0);
        run<16>("4 addresses, one per quarter-warp", 3, nw, out, cyc, sms, 0);
        run<16>("2 addresses, one per half-warp", 4, nw, out, cyc, sms, 0);
        run<16>("32 rows at a 304-byte pitch", 5, nw, out, cyc, sms, 1);
        run<16>("32 rows at a 2416-byte pitch", 6, nw, out, cyc, sms, 1);
        run<8>("32 distinct addresses (256 B)", 0, nw, out, cyc, sms, 0);
        run<8>("one address for the warp", 1, nw, out, cyc, sms, 0);
        run<8>("16 addresses, the same in each half-warp", 7, nw, out, cyc, sms, 0);
        run<8>("2 addresses, one per half-warp", 4, nw, out, cyc, sms, 0);
        run<4>("32 distinct addresses (128 B)", 0, nw, out, cyc, sms, 0);
        run<4>("one address for the warp", 1, nw, out, cyc, sms, 0);
    }
    return 0;
}
