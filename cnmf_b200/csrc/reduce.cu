// K2 -- cross-CTA reduction of the H-update partials and the H finalise step.
//
//   reduce_partials : red[e] = sum over partial blocks (fixed order, fp64);
//                     red[n_entries] = sum of the per-warp chi2 partials.
//   (multi-GPU: ncclAllReduce(red) sits between the two kernels)
//   finalize        : hist[idx] = chi2; tolerance test; H <- H * num / max(den, eps)
//
// Both are tiny (k*ld*2 + 1 doubles out); they exist so that the k x M sums are
// deterministic and in fp64 whatever the compute dtype.
#include "common.cuh"
#include "launch.h"

namespace cnmf {

// grid = ceil(n_entries / 32) + 1, block = 256: thread (e = tid & 31, slice = tid >> 5)
__global__ void __launch_bounds__(256) reduce_partials_kernel(const double *__restrict__ part, int nblk,
                                                              long long stride, int n_entries,
                                                              const double *__restrict__ chi_part, int n_chi,
                                                              double *__restrict__ red, const State *state) {
    if (state->done) return;
    __shared__ double sh[256];
    const int tid = threadIdx.x;
    if (blockIdx.x == gridDim.x - 1) {
        // chi2: strided sum then a fixed-order tree
        double c = 0.0;
        for (int i = tid; i < n_chi; i += 256) c += chi_part[i];
        sh[tid] = c;
        __syncthreads();
        for (int o = 128; o > 0; o >>= 1) {
            if (tid < o) sh[tid] += sh[tid + o];
            __syncthreads();
        }
        if (tid == 0) red[n_entries] = sh[0];
        return;
    }
    const int e = blockIdx.x * 32 + (tid & 31);
    const int slice = tid >> 5;
    double a0 = 0.0, a1 = 0.0;
    if (e < n_entries) {
        int b = slice;
        for (; b + 8 < nblk; b += 16) {
            a0 += part[(long long)b * stride + e];
            a1 += part[(long long)(b + 8) * stride + e];
        }
        if (b < nblk) a0 += part[(long long)b * stride + e];
    }
    sh[tid] = a0 + a1;
    __syncthreads();
    if (slice == 0 && e < n_entries) {
        double s = sh[tid];
#pragma unroll
        for (int k = 1; k < 8; ++k) s += sh[tid + 32 * k];
        red[e] = s;
    }
}

// grid = ceil(k * m / 256)
template <typename T>
__global__ void __launch_bounds__(256) finalize_kernel(const double *__restrict__ red, T *__restrict__ H, State *state,
                                                       double *__restrict__ hist, int k, int m, long long ld, double tol,
                                                       T eps, int update_h) {
    if (state->done) return;
    const int idx = state->k1_count - 1;
    const int n_entries = 2 * k * (int)ld;
    const double chi = red[n_entries];
    bool conv = false;
    if (tol > 0.0 && idx >= 1) {
        const double prev = hist[idx - 1];
        conv = (prev - chi) <= tol * prev;
    }
    const bool bad = !(chi == chi) || chi > 1.7e308 || chi < -1.7e308;
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (update_h && !conv && !bad && i < k * m) {
        const int a = i / m, j = i - a * m;
        const T num = (T)red[(long long)a * ld + j];
        const T den = (T)red[(long long)k * ld + (long long)a * ld + j];
        H[i] = (H[i] * num) / (den > eps ? den : eps);
    }
    if (i == 0) {
        // hist[idx] is read (as prev) only by the NEXT finalize launch
        hist[idx] = chi;
        if (bad) {
            state->nonfinite = 1;
            state->done = 1;
            state->n_iter = idx;
        } else if (conv) {
            state->done = 1;
            state->n_iter = idx;
        }
    }
}

void launch_reduce(const double *part, int nblk, long long stride, int n_entries, const double *chi_part, int n_chi,
                   double *red, const State *state, cudaStream_t st) {
    const int grid = (n_entries + 31) / 32 + 1;
    reduce_partials_kernel<<<grid, 256, 0, st>>>(part, nblk, stride, n_entries, chi_part, n_chi, red, state);
    count_launch();
}

void launch_finalize(int dtype, const double *red, void *H, State *state, double *hist, int k, int m, long long ld,
                     double tol, double eps, int update_h, cudaStream_t st) {
    const int grid = (k * m + 255) / 256;
    if (dtype == 0)
        finalize_kernel<float><<<grid, 256, 0, st>>>(red, (float *)H, state, hist, k, m, ld, tol, (float)eps, update_h);
    else
        finalize_kernel<double><<<grid, 256, 0, st>>>(red, (double *)H, state, hist, k, m, ld, tol, eps, update_h);
    count_launch();
}

}  // namespace cnmf
