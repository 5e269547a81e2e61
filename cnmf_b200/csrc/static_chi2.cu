// cnmf_chi2_static: chi2 = sum V (X - WH)^2 of a given (W, H) for callers that may not allocate
// (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  A variant of the
// chi2-only pass (K3, cnmf_chi2) whose scratch -- the per-row-block partial sums and the arrival
// counter of the last-block reduction -- lives in ONE file-scope device array of a fixed size
// instead of a plan's cudaMalloc'ed block.  No counterpart in the
// reference.
//
// * Every GPU that loads the library holds the array (262 400 bytes) until the process exits.
// * Every call shares it: calls that overlap on one GPU (two streams, two host threads) race.
// * The array keeps the last call's contents.  The partials a call reads are all written by that
//   call first; the arrival counter must start at zero and is cleared on the stream before each
//   launch (and reset by the last block, so an aborted call cannot poison the next one).
// * In host code `g_chi2_scratch` is a host-side stand-in, not a device address: the kernel gets
//   the pointer cudaGetSymbolAddress returned for the CURRENT device (looked up once per device).
#include <cuda_runtime.h>
#include <mutex>

#include "../../include/cnmf_b200.h"
#include "launch.h"

namespace cnmf {
void set_error(const char *fmt, ...);

namespace {

constexpr int kRowsPerBlock = 1024;  // one partial per block of rows: the sum order depends on n_rows only
constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kMaxBlocks = CNMF_CHI2_STATIC_MAX_ROWS / kRowsPerBlock;  // 32 768 partials
constexpr int kHeader = 32;                                            // 256 bytes: [0] = arrival counter

static_assert(CNMF_CHI2_STATIC_MAX_ROWS % kRowsPerBlock == 0, "row limit must be a whole number of blocks");

static __device__ __align__(256) double g_chi2_scratch[kHeader + kMaxBlocks];

template <typename T>
__global__ void __launch_bounds__(kThreads) chi2_static_kernel(const T *__restrict__ X, const T *__restrict__ V,
                                                               const T *__restrict__ W, const T *__restrict__ H,
                                                               long long n_rows, long long ld, int m, int k,
                                                               double *scratch, double *out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T *Hs = reinterpret_cast<T *>(smem_raw);
    __shared__ double wsum[kWarps];
    __shared__ double tsum[kThreads];
    __shared__ int is_last;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < k * m; i += kThreads) Hs[i] = H[i];
    __syncthreads();

    const long long r0 = (long long)blockIdx.x * kRowsPerBlock;
    const long long r1 = r0 + kRowsPerBlock < n_rows ? r0 + kRowsPerBlock : n_rows;
    double acc = 0.0;
    for (long long row = r0 + warp; row < r1; row += kWarps) {
        T w[CNMF_MAX_K];
#pragma unroll
        for (int a = 0; a < CNMF_MAX_K; ++a) w[a] = a < k ? W[row * k + a] : T(0);
        const T *xr = X + row * ld, *vr = V + row * ld;
        for (int j = lane; j < m; j += 32) {
            T wh = T(0);
#pragma unroll
            for (int a = 0; a < CNMF_MAX_K; ++a)
                if (a < k) wh = fma(w[a], Hs[a * m + j], wh);
            const T d = xr[j] - wh;
            acc += (double)(vr[j] * d * d);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) wsum[warp] = acc;
    __syncthreads();
    double *part = scratch + kHeader;
    unsigned *arrived = reinterpret_cast<unsigned *>(scratch);
    if (tid == 0) {
        double s = 0.0;
        for (int i = 0; i < kWarps; ++i) s += wsum[i];
        part[blockIdx.x] = s;
        __threadfence();
        is_last = atomicAdd(arrived, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!is_last) return;
    // last block to arrive: fixed-order sum of all partials (thread t takes t, t + 256, ...; then a tree)
    __threadfence();
    double s = 0.0;
    for (int i = tid; i < (int)gridDim.x; i += kThreads) s += __ldcg(part + i);
    tsum[tid] = s;
    __syncthreads();
    for (int o = kThreads / 2; o > 0; o >>= 1) {
        if (tid < o) tsum[tid] += tsum[tid + o];
        __syncthreads();
    }
    if (tid == 0) {
        *out = tsum[0];
        *arrived = 0u;  // the next call finds the counter at zero even without its own clear
    }
}

std::mutex g_mu;
double *g_scratch_ptr[kMaxDevices] = {};

// device address of the scratch array on the current device (looked up once per device)
int scratch_for_current_device(double **ptr) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        set_error("cnmf_chi2_static: cudaGetDevice: CUDA error %d (%s)", (int)e, cudaGetErrorString(e));
        return 1000 + (int)e;
    }
    if (dev < 0 || dev >= kMaxDevices) {
        set_error("cnmf_chi2_static: device ordinal %d is beyond the %d this library keeps state for", dev, kMaxDevices);
        return CNMF_ERR_UNSUPPORTED;
    }
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_ptr[dev]) {
        double *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_chi2_scratch);
        if (e != cudaSuccess || !p) {
            set_error("cnmf_chi2_static: cudaGetSymbolAddress on device %d: CUDA error %d (%s)", dev, (int)e,
                      cudaGetErrorString(e));
            return e != cudaSuccess ? 1000 + (int)e : CNMF_ERR_NODEVICE;
        }
        g_scratch_ptr[dev] = p;
    }
    *ptr = g_scratch_ptr[dev];
    return 0;
}

}  // namespace
}  // namespace cnmf

extern "C" int cnmf_chi2_static(int32_t dtype, const void *dX, const void *dV, const void *dW, const void *dH,
                                int64_t n_rows, int64_t ld, int32_t m, int32_t k, double *d_out, void *stream) {
    using namespace cnmf;
    if (!dX || !dV || !dW || !dH || !d_out || (dtype != CNMF_F32 && dtype != CNMF_F64) || k < 1 || k > CNMF_MAX_K ||
        m < 1 || ld < m || n_rows < 1) {
        set_error("cnmf_chi2_static: bad arguments (NULL pointer, dtype, k outside 1..%d, m < 1, ld < m or n_rows < 1)",
                  CNMF_MAX_K);
        return CNMF_ERR_ARG;
    }
    if (n_rows > (int64_t)CNMF_CHI2_STATIC_MAX_ROWS) {
        set_error("cnmf_chi2_static: n_rows = %lld exceeds CNMF_CHI2_STATIC_MAX_ROWS = %d (the static scratch holds one "
                  "partial sum per %d rows); use cnmf_chi2 on a plan",
                  (long long)n_rows, CNMF_CHI2_STATIC_MAX_ROWS, kRowsPerBlock);
        return CNMF_ERR_UNSUPPORTED;
    }
    const size_t h_bytes = (size_t)k * (size_t)m * (dtype == CNMF_F64 ? 8 : 4);
    if (h_bytes > (size_t)CNMF_CHI2_STATIC_MAX_H_BYTES) {
        set_error("cnmf_chi2_static: H needs %zu bytes of shared memory, the limit is CNMF_CHI2_STATIC_MAX_H_BYTES = %d; "
                  "use cnmf_chi2 on a plan",
                  h_bytes, CNMF_CHI2_STATIC_MAX_H_BYTES);
        return CNMF_ERR_UNSUPPORTED;
    }
    double *scratch = nullptr;
    if (int e = scratch_for_current_device(&scratch)) return e;
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(scratch, 0, kHeader * sizeof(double), st);  // the arrival counter starts at zero
    if (e != cudaSuccess) {
        set_error("cnmf_chi2_static: cudaMemsetAsync: CUDA error %d (%s)", (int)e, cudaGetErrorString(e));
        return 1000 + (int)e;
    }
    const int grid = (int)((n_rows + kRowsPerBlock - 1) / kRowsPerBlock);
    if (dtype == CNMF_F64)
        chi2_static_kernel<double><<<grid, kThreads, h_bytes, st>>>((const double *)dX, (const double *)dV,
                                                                    (const double *)dW, (const double *)dH, n_rows, ld,
                                                                    m, k, scratch, d_out);
    else
        chi2_static_kernel<float><<<grid, kThreads, h_bytes, st>>>((const float *)dX, (const float *)dV,
                                                                   (const float *)dW, (const float *)dH, n_rows, ld, m,
                                                                   k, scratch, d_out);
    count_launch();
    e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("cnmf_chi2_static: launch: CUDA error %d (%s)", (int)e, cudaGetErrorString(e));
        return 1000 + (int)e;
    }
    return 0;
}
