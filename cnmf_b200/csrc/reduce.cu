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

// ---------------------------------------------------------------------------------------------
// Row-sharded mode with peer windows (comm.h): the all-reduce of red[] and the finalise step in
// ONE single-CTA kernel.  Each rank stores its partial sums into slot[seq & 1][rank] of every
// rank's window (NVLink stores, fire and forget), fences, then releases flag[seq & 1][rank] =
// seq in every window; it then waits on the flags of its OWN window (local polling), sums the
// world slots in rank order -- the same order everywhere, so H and chi2 stay bit-identical
// across ranks -- and applies the H update.  Two slots suffice: a rank can write exchange
// seq + 2 into a slot only after it saw every peer's flag seq + 1, which a peer sends only after
// it finished reading exchange seq.  The wait is bounded (PeerView::timeout_ns, 120 s unless
// cnmf_comm_set_timeout changed it: a dead peer ends the solve with State::peer_timeout instead of
// hanging the GPU; after a timeout the communicator is out of step and must be destroyed).
__device__ __forceinline__ void st_release_sys(unsigned int *p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int *p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ double ld_volatile(const double *p) {
    double v;
    asm volatile("ld.volatile.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

constexpr int kExchangeThreads = 1024;

// One rank's exchange + finalise, executed by one CTA.
template <typename T>
__device__ __forceinline__ void exchange_body(const PeerView &pv, double *__restrict__ red, int count, T *__restrict__ H,
                                              State *state, double *__restrict__ hist, int k, int m, long long ld, double tol,
                                              T eps, int update_h, int finalize) {
    if (state->done) return;
    __shared__ unsigned int s_seq;
    __shared__ int s_fail;
    const int tid = threadIdx.x;
    if (tid == 0) {
        s_seq = *pv.seq + 1;
        s_fail = 0;
    }
    __syncthreads();
    const unsigned int seq = s_seq;
    const int par = (int)(seq & 1u);
    // push this rank's partial sums to every window (own included)
    for (int e = tid; e < count; e += kExchangeThreads) {
        const double v = red[e];
#pragma unroll 4
        for (int r = 0; r < pv.world; ++r) peer_slot(pv.win[r], par, pv.world, pv.rank)[e] = v;
    }
    __threadfence_system();
    __syncthreads();
    if (tid < pv.world) {
        st_release_sys(&pv.win[tid]->flag[par][pv.rank], seq);
        const unsigned int *f = &pv.win[pv.rank]->flag[par][tid];
        const unsigned long long t0 = global_ns();
        while (ld_acquire_sys(f) != seq) {
            if (global_ns() - t0 > pv.timeout_ns) {
                s_fail = 1;
                break;
            }
        }
    }
    __syncthreads();
    if (s_fail) {
        if (tid == 0) {
            state->peer_timeout = 1;
            state->done = 1;
            state->n_iter = state->k1_count > 0 ? state->k1_count - 1 : 0;
        }
        return;
    }
    PeerWindow *mine = pv.win[pv.rank];
    for (int e = tid; e < count; e += kExchangeThreads) {
        double s = 0.0;
        for (int r = 0; r < pv.world; ++r) s += ld_volatile(peer_slot(mine, par, pv.world, r) + e);
        red[e] = s;
    }
    if (tid == 0) *pv.seq = seq;
    __syncthreads();
    if (!finalize) return;
    // as finalize_kernel, one CTA
    const int idx = state->k1_count - 1;
    const int n_entries = 2 * k * (int)ld;
    const double chi = red[n_entries];
    bool conv = false;
    if (tol > 0.0 && idx >= 1) {
        const double prev = hist[idx - 1];
        conv = (prev - chi) <= tol * prev;
    }
    const bool bad = !(chi == chi) || chi > 1.7e308 || chi < -1.7e308;
    if (update_h && !conv && !bad)
        for (int i = tid; i < k * m; i += kExchangeThreads) {
            const int a = i / m, j = i - a * m;
            const T num = (T)red[(long long)a * ld + j];
            const T den = (T)red[(long long)k * ld + (long long)a * ld + j];
            H[i] = (H[i] * num) / (den > eps ? den : eps);
        }
    if (tid == 0) {
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

template <typename T>
__global__ void __launch_bounds__(kExchangeThreads) exchange_finalize_kernel(PeerView pv, double *__restrict__ red, int count,
                                                                             T *__restrict__ H, State *state,
                                                                             double *__restrict__ hist, int k, int m,
                                                                             long long ld, double tol, T eps, int update_h,
                                                                             int finalize) {
    exchange_body<T>(pv, red, count, H, state, hist, k, m, ld, tol, eps, update_h, finalize);
}

// Several ranks of a single-process group that live on ONE device (virtual shards, tests): their
// exchanges run as ONE cooperative launch, CTA r doing rank r's exchange, so the CTAs that wait on each
// other's flags are co-resident by construction.  (Separate launches that spin on one another have no such
// guarantee on one GPU.)  The per-CTA code is exchange_body, the same as in the production kernel.
template <typename T>
__global__ void __launch_bounds__(kExchangeThreads) exchange_group_kernel(const GroupRank *__restrict__ ranks, int count, int k,
                                                                          int m, long long ld, double tol, T eps,
                                                                          int update_h, int finalize) {
    const GroupRank g = ranks[blockIdx.x];
    exchange_body<T>(g.pv, g.red, count, (T *)g.H, g.state, g.hist, k, m, ld, tol, eps, update_h, finalize);
}

int launch_exchange_group(int dtype, const GroupRank *d_ranks, int world, int count, int k, int m, long long ld, double tol,
                          double eps, int update_h, int finalize, cudaStream_t st) {
    float epsf = (float)eps;
    void *args_f[] = {(void *)&d_ranks, &count, &k, &m, &ld, &tol, &epsf, &update_h, &finalize};
    void *args_d[] = {(void *)&d_ranks, &count, &k, &m, &ld, &tol, &eps, &update_h, &finalize};
    cudaError_t e = dtype == 0 ? cudaLaunchCooperativeKernel((const void *)exchange_group_kernel<float>, dim3(world),
                                                             dim3(kExchangeThreads), args_f, 0, st)
                               : cudaLaunchCooperativeKernel((const void *)exchange_group_kernel<double>, dim3(world),
                                                             dim3(kExchangeThreads), args_d, 0, st);
    count_launch();
    return (int)e;
}

void launch_exchange(int dtype, const PeerView &pv, double *red, int count, void *H, State *state, double *hist, int k,
                     int m, long long ld, double tol, double eps, int update_h, int finalize, cudaStream_t st) {
    if (dtype == 0)
        exchange_finalize_kernel<float><<<1, kExchangeThreads, 0, st>>>(pv, red, count, (float *)H, state, hist, k, m, ld, tol,
                                                                        (float)eps, update_h, finalize);
    else
        exchange_finalize_kernel<double><<<1, kExchangeThreads, 0, st>>>(pv, red, count, (double *)H, state, hist, k, m, ld,
                                                                         tol, eps, update_h, finalize);
    count_launch();
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
