// Operand preparation (pad / mask-fold / validate) and the device generator of
// synthetic spectra (K5).  Both run once per solve, not per iteration.
#include "common.cuh"
#include "launch.h"
#include "../../include/cnmf_b200.h"

namespace cnmf {

template <typename T>
__global__ void __launch_bounds__(256) prepare_kernel(const T *__restrict__ sx, const T *__restrict__ sw,
                                                      const uint8_t *__restrict__ mk, long long src_ld,
                                                      long long mask_ld, T *dx, T *dv, long long dst_ld,
                                                      long long n_rows, int m, int *status) {
    const long long total = n_rows * dst_ld;
    int bad = 0;
    for (long long idx = (long long)blockIdx.x * 256 + threadIdx.x; idx < total; idx += (long long)gridDim.x * 256) {
        const long long i = idx / dst_ld;
        const int j = (int)(idx - i * dst_ld);
        T x = T(0), v = T(0);
        if (j < m) {
            x = sx[i * src_ld + j];
            T w = sw ? sw[i * src_ld + j] : T(1);
            if (!isfinite(w)) {
                bad |= CNMF_BAD_W_NONFINITE;
                w = T(0);
            } else if (w < T(0)) {
                bad |= CNMF_BAD_W_NEGATIVE;
                w = T(0);
            }
            const bool good = mk ? (mk[i * mask_ld + j] != 0) : true;
            v = good ? w : T(0);
            if (v > T(0)) {
                if (!isfinite(x)) {
                    bad |= CNMF_BAD_X_NONFINITE;
                    x = T(0);
                    v = T(0);
                } else if (x < T(0)) {
                    bad |= CNMF_BAD_X_NEGATIVE;
                }
            } else {
                x = T(0);
            }
        }
        dx[idx] = x;
        dv[idx] = v;
    }
    if (bad && status) atomicOr(status, bad);
}

int launch_prepare(int dtype, const void *src_x, const void *src_w, const uint8_t *src_mask, long long src_ld,
                   long long mask_ld, void *dst_x, void *dst_v, long long dst_ld, long long n_rows, int m,
                   int *d_status, cudaStream_t st) {
    if (n_rows <= 0) return 0;
    const long long total = n_rows * dst_ld;
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 32) blocks = 148LL * 32;
    if (dtype == CNMF_F32)
        prepare_kernel<float><<<(int)blocks, 256, 0, st>>>((const float *)src_x, (const float *)src_w, src_mask, src_ld,
                                                           mask_ld, (float *)dst_x, (float *)dst_v, dst_ld, n_rows, m,
                                                           d_status);
    else
        prepare_kernel<double><<<(int)blocks, 256, 0, st>>>((const double *)src_x, (const double *)src_w, src_mask,
                                                            src_ld, mask_ld, (double *)dst_x, (double *)dst_v, dst_ld,
                                                            n_rows, m, d_status);
    count_launch();
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------
// synthetic spectra: same hash and formulas as cnmf_b200/synth.py
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint64_t synth_hash(uint64_t seed, uint64_t stream, uint64_t idx) {
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (idx + 1ull) + 0xD1B54A32D192ED03ull * stream;
    z ^= z >> 30;
    z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27;
    z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}
__device__ __forceinline__ void synth_uniforms(uint64_t h, double &u1, double &u2) {
    u1 = ((double)(uint32_t)(h >> 32) + 0.5) * (1.0 / 4294967296.0);
    u2 = ((double)(uint32_t)(h & 0xFFFFFFFFull) + 0.5) * (1.0 / 4294967296.0);
}
__device__ __forceinline__ double synth_normal(uint64_t seed, uint64_t stream, uint64_t idx) {
    double u1, u2;
    synth_uniforms(synth_hash(seed, stream, idx), u1, u2);
    return sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
}

constexpr int kSynthRows = 8;

template <typename T>
__global__ void __launch_bounds__(256) synth_kernel(T *X, T *V, long long n_rows, long long row0, int m, long long ld,
                                                    int k, const double *__restrict__ hstar, uint64_t seed, int vmode,
                                                    double p_missing) {
    __shared__ double wstar[kSynthRows * CNMF_MAX_K];
    const long long r_base = (long long)blockIdx.x * kSynthRows;
    const int rows = (int)((n_rows - r_base) < kSynthRows ? (n_rows - r_base) : kSynthRows);
    if (threadIdx.x < rows * k) {
        const int r = threadIdx.x / k, a = threadIdx.x - r * k;
        const uint64_t grow = (uint64_t)(row0 + r_base + r);
        wstar[r * CNMF_MAX_K + a] = exp(0.5 * synth_normal(seed, 1, grow * (uint64_t)k + a));
    }
    __syncthreads();
    const int total = rows * (int)ld;
    for (int idx = threadIdx.x; idx < total; idx += 256) {
        const int r = idx / (int)ld, j = idx - r * (int)ld;
        double x = 0.0, v = 0.0;
        if (j < m) {
            double s = 0.0;
            for (int a = 0; a < k; ++a) s += wstar[r * CNMF_MAX_K + a] * hstar[a * m + j];
            const double sig = 0.02 * s + 1e-4;
            const uint64_t e = (uint64_t)(row0 + r_base + r) * (uint64_t)m + (uint64_t)j;
            x = fmax(s + sig * synth_normal(seed, 2, e), 0.0);
            v = (vmode & 1) ? 1.0 / (sig * sig) : 1.0;
            if (vmode & 2) {
                double u1, u2;
                synth_uniforms(synth_hash(seed, 3, e), u1, u2);
                if (u1 < p_missing) {
                    x = 0.0;
                    v = 0.0;
                }
            }
        }
        X[(r_base + r) * ld + j] = (T)x;
        V[(r_base + r) * ld + j] = (T)v;
    }
}

int launch_synth(int dtype, void *dX, void *dV, long long n_rows, long long row0, int m, long long ld, int k,
                 const double *d_hstar, uint64_t seed, int vmode, double p_missing, cudaStream_t st) {
    if (n_rows <= 0) return 0;
    const long long blocks = (n_rows + kSynthRows - 1) / kSynthRows;
    if (blocks > 0x7fffffffLL) return (int)cudaErrorInvalidValue;
    if (dtype == CNMF_F32)
        synth_kernel<float><<<(int)blocks, 256, 0, st>>>((float *)dX, (float *)dV, n_rows, row0, m, ld, k, d_hstar,
                                                         seed, vmode, p_missing);
    else
        synth_kernel<double><<<(int)blocks, 256, 0, st>>>((double *)dX, (double *)dV, n_rows, row0, m, ld, k, d_hstar,
                                                          seed, vmode, p_missing);
    count_launch();
    return (int)cudaGetLastError();
}

}  // namespace cnmf
