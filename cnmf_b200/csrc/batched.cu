// K4 -- batched small-problem solver: one CTA per independent problem, all
// iterations inside the kernel (config 4: 4096 x (512 x 81), k = 3).
//
// H lives in shared memory for the whole solve; X and V tiles are streamed
// through a cp.async double buffer (after the first pass they come from L2: a
// wave of resident CTAs re-reads far less than the 126 MB L2 holds), W is
// updated in place in global memory.  Per tile: phase 1 (G lanes per row ->
// W') then phase 2 (thread owns one pack of columns -> chi2 and the k x M
// partials in registers).  After the last tile the column streams are combined
// in fp64 through shared memory, H is updated, chi2 recorded and the
// tolerance test applied -- no inter-CTA communication at all.
#include <stdio.h>
#include <stdlib.h>

#include "phases.cuh"
#include "launch.h"

namespace cnmf {

constexpr int kBT = 256;   // threads per problem
// BG lanes per row in phase 1 -> BR = 256 / BG rows per tile; OCC = CTAs per SM the kernel is
// compiled for (register budget).  (4, 2) exists for every k; k <= 4 adds the smaller tiles.
__host__ __device__ constexpr int batched_rows(int BG) { return kBT / BG; }

template <typename T> struct BatchedParams {
    const T *X;
    const T *V;
    T *W;
    T *H;
    double *hist;
    int *niter;
    long long n_rows, ld;
    int m, max_iter, update;
    int wpre;  // W rows of a tile come with its cp.async group (else: global loads at the head of phase 1)
    double tol;
    T eps;
};

struct BatchedSmem {
    int h_off, acc_off, wd_off, win_off, win_bytes, buf_off, buf_bytes, red_off, total;
};
__host__ __device__ inline BatchedSmem batched_smem(int K, long long ld, int sz, int kBR) {
    BatchedSmem s;
    auto up = [](long long v) { return (int)((v + 127) / 128 * 128); };
    const int KP = (K + 1) & ~1;
    s.h_off = 0;
    s.acc_off = up((long long)K * ld * sz);
    s.wd_off = s.acc_off + up(2LL * K * ld * 8);
    s.red_off = s.wd_off + up((long long)kBR * KP * 8);
    s.win_off = s.red_off + 128;  // W rows of the tile in flight, two buffers
    s.win_bytes = up((long long)kBR * K * sz);
    s.buf_off = s.win_off + 2 * s.win_bytes;
    s.buf_bytes = up((long long)kBR * ld * sz);
    s.total = s.buf_off + 4 * s.buf_bytes;  // {X, V} x 2 buffers
    return s;
}

__device__ __forceinline__ void cp_async16(void *dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_elem(float *dst, const float *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_elem(double *dst, const double *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
template <int N> __device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <typename T, int K, int kBG, int OCC>
__global__ void __launch_bounds__(kBT, OCC) batched_kernel(const BatchedParams<T> p) {
    constexpr int kBR = batched_rows(kBG);
    typedef Pk<T> O;
    typedef typename O::P P;
    constexpr int SZ = sizeof(T);
    constexpr int PW = O::W;
    constexpr int KP = (K + 1) & ~1;
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long ld = p.ld;
    const int ldv = (int)(ld * SZ / 16);
    const int packs = (int)(ld / PW);
    const BatchedSmem L = batched_smem(K, ld, SZ, kBR);
    T *Hs = reinterpret_cast<T *>(smem + L.h_off);
    double *acc = reinterpret_cast<double *>(smem + L.acc_off);  // [2][K][ld]
    P *Wd = reinterpret_cast<P *>(smem + L.wd_off);
    double *red = reinterpret_cast<double *>(smem + L.red_off);  // [8] warp sums, [8] chi2 now, [9] flag
    const long long b = blockIdx.x;
    const T *X = p.X + b * p.n_rows * ld;
    const T *V = p.V + b * p.n_rows * ld;
    T *W = p.W + b * p.n_rows * K;
    T *Hg = p.H + b * (long long)K * p.m;
    double *hist = p.hist + b * (long long)(p.max_iter + 1);
    const int n_tiles = (int)((p.n_rows + kBR - 1) / kBR);
    const int streams = kBT / packs;  // >= 1 (checked on the host)
    const int stream = tid / packs, ci = tid - stream * packs;
    const bool active = stream < streams;
    const int r1 = tid / kBG, sub = tid % kBG;

    for (int i = tid; i < K * (int)ld; i += kBT) {
        const int a = i / (int)ld, j = i - a * (int)ld;
        Hs[i] = (j < p.m) ? Hg[(long long)a * p.m + j] : T(0);
    }
    for (int i = tid; i <= p.max_iter; i += kBT) hist[i] = __longlong_as_double(0x7ff8000000000000LL);
    __syncthreads();

    auto load_tile = [&](int t, int buf) {
        const long long r0 = (long long)t * kBR;
        const int rows = (int)((p.n_rows - r0) < kBR ? (p.n_rows - r0) : kBR);
        const int nvec = rows * ldv;
        unsigned char *xb = smem + L.buf_off + (2 * buf) * L.buf_bytes;
        unsigned char *vb = xb + L.buf_bytes;
        const unsigned char *xs = reinterpret_cast<const unsigned char *>(X + r0 * ld);
        const unsigned char *vs = reinterpret_cast<const unsigned char *>(V + r0 * ld);
        for (int i = tid; i < nvec; i += kBT) {
            cp_async16(xb + i * 16, xs + (long long)i * 16);
            cp_async16(vb + i * 16, vs + (long long)i * 16);
        }
        // the tile's W rows ride along (they were read from global memory at the head of phase 1,
        // an exposed L2 round trip per tile); rows of a later tile are never written before it
        T *wb = reinterpret_cast<T *>(smem + L.win_off + buf * L.win_bytes);
        if (p.wpre)
            for (int i = tid; i < rows * K; i += kBT) cp_async_elem(wb + i, W + r0 * K + i);
        cp_async_commit();
    };

    double prev_chi = 0.0;
    int it_done = 0;
    // pass 0 evaluates (W0, H0); pass it >= 1 first applies the H update from
    // the previous pass's partials, then updates W tile by tile
    for (int pass = 0; pass <= p.max_iter; ++pass) {
        const bool upd_w = pass > 0 && (p.update & 1);
        P Hr[K], nH[K], dH[K];
#pragma unroll
        for (int a = 0; a < K; ++a) {
            Hr[a] = active ? *reinterpret_cast<const P *>(reinterpret_cast<const unsigned char *>(Hs) +
                                                         ((long long)a * ld + ci * PW) * SZ)
                           : O::zero();
            nH[a] = O::zero();
            dH[a] = O::zero();
        }
        double chid = 0.0;
        load_tile(0, 0);
        for (int t = 0; t < n_tiles; ++t) {
            const int buf = t & 1;
            if (t + 1 < n_tiles) {
                load_tile(t + 1, buf ^ 1);
                cp_async_wait<1>();
            } else {
                cp_async_wait<0>();
            }
            __syncthreads();
            const long long r0 = (long long)t * kBR;
            const int rows = (int)((p.n_rows - r0) < kBR ? (p.n_rows - r0) : kBR);
            const unsigned char *xb = smem + L.buf_off + (2 * buf) * L.buf_bytes;
            const unsigned char *vb = xb + L.buf_bytes;
            {   // phase 1
                const bool valid = r1 < rows;
                const int rr = valid ? r1 : 0;
                T w[K], wn[K];
                const T *wb = reinterpret_cast<const T *>(smem + L.win_off + buf * L.win_bytes);
#pragma unroll
                for (int a = 0; a < K; ++a) w[a] = p.wpre ? wb[rr * K + a] : W[(r0 + rr) * K + a];
                if (upd_w) {
                    phase1_row<T, K, kBG>(xb + (long long)rr * ld * SZ, vb + (long long)rr * ld * SZ,
                                          reinterpret_cast<const unsigned char *>(Hs), ldv, sub, w, p.eps, wn);
                } else {
#pragma unroll
                    for (int a = 0; a < K; ++a) wn[a] = w[a];
                }
                if (valid) {
#pragma unroll
                    for (int a = 0; a < K; ++a)
                        if ((a % kBG) == sub) {
                            if (upd_w) W[(r0 + rr) * K + a] = wn[a];
                            Wd[rr * KP + a] = O::dup(wn[a]);
                        }
                }
            }
            __syncthreads();
            if (active) {   // phase 2
                P chi = O::zero();
                for (int r = stream; r < rows; r += streams) {
                    const P x = *reinterpret_cast<const P *>(xb + ((long long)r * ld + ci * PW) * SZ);
                    const P v = *reinterpret_cast<const P *>(vb + ((long long)r * ld + ci * PW) * SZ);
                    P wd[K];
#pragma unroll
                    for (int a = 0; a < K; ++a) wd[a] = Wd[r * KP + a];
                    phase2_pack<T, K>(x, v, wd, Hr, nH, dH, chi);
                }
                chid += O::dsum(chi);
            }
            __syncthreads();
        }
        // ---- combine the column streams in fp64, chi2, tolerance, H update
        for (int i = tid; i < 2 * K * (int)ld; i += kBT) acc[i] = 0.0;
        __syncthreads();
        for (int s = 0; s < streams; ++s) {
            if (active && stream == s) {
#pragma unroll
                for (int a = 0; a < K; ++a)
#pragma unroll
                    for (int e = 0; e < PW; ++e) {
                        const int idx = a * (int)ld + ci * PW + e;
                        acc[idx] += (double)O::lane(nH[a], e);
                        acc[K * (int)ld + idx] += (double)O::lane(dH[a], e);
                    }
            }
            __syncthreads();
        }
        chid = warp_sum(chid);
        if (lane == 0) red[warp] = chid;
        __syncthreads();
        if (tid == 0) {
            double c = 0.0;
            for (int w = 0; w < kBT / 32; ++w) c += red[w];
            hist[pass] = c;
            int stop = 0;
            const bool bad = !(c == c) || c > 1.7e308;
            if (bad) stop = 1;
            if (p.tol > 0.0 && pass >= 1 && (prev_chi - c) <= p.tol * prev_chi) stop = 1;
            red[8] = c;
            reinterpret_cast<int *>(red + 9)[0] = stop;
        }
        __syncthreads();
        prev_chi = red[8];
        it_done = pass;
        if (reinterpret_cast<int *>(red + 9)[0] || pass == p.max_iter) break;
        if (p.update & 2) {
            for (int i = tid; i < K * p.m; i += kBT) {
                const int a = i / p.m, j = i - a * p.m;
                const T num = (T)acc[a * (int)ld + j];
                const T den = (T)acc[K * (int)ld + a * (int)ld + j];
                const T h = Hs[a * (int)ld + j];
                Hs[a * (int)ld + j] = (h * num) / (den > p.eps ? den : p.eps);
            }
        }
        __syncthreads();
    }
    for (int i = tid; i < K * p.m; i += kBT) {
        const int a = i / p.m, j = i - a * p.m;
        Hg[i] = Hs[a * (int)ld + j];
    }
    if (tid == 0) p.niter[b] = it_done;
}

static void batched_variant(int k, int dtype, int *bg, int *occ) {
    // fp64 rows are 2x the vectors: 8 lanes per row (32-row tiles) measured 14 % faster at config 4
    int g = (dtype == 1 && k <= 4) ? 8 : 4, o = 2;
    if (k <= 4) {
        const char *e = getenv("CNMF_BATCHED");  // "BG,OCC" -- sweeps only
        int eg = 0, eo = 0;
        if (e && sscanf(e, "%d,%d", &eg, &eo) == 2 && (eg == 4 || eg == 8 || eg == 16) && (eo == 2 || eo == 3 || eo == 4)) {
            g = eg;
            o = eo;
        }
    }
    *bg = g;
    *occ = o;
}

int batched_smem_bytes(int dtype, int k, long long ld, int *threads_out) {
    const int sz = dtype == 0 ? 4 : 8, pw = dtype == 0 ? 2 : 1;
    if (threads_out) *threads_out = kBT;
    if (ld / pw > kBT) return -1;  // one pack of columns per thread
    int bg, occ;
    batched_variant(k, dtype, &bg, &occ);
    return batched_smem(k, ld, sz, batched_rows(bg)).total;
}

template <typename T, int K, int BG, int OCC> static int launch_bv(const BatchedArgs &a, cudaStream_t st) {
    auto kern = batched_kernel<T, K, BG, OCC>;
    const int smem = batched_smem(K, a.ld, sizeof(T), batched_rows(BG)).total;
    static int smem_set[kMaxDevices] = {0};  // the opt-in above 48 KB is a per-device function attribute
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= kMaxDevices) return (int)cudaErrorInvalidDevice;
    if (smem > smem_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return (int)e;
        smem_set[dev] = smem;
    }
    BatchedParams<T> p;
    p.X = (const T *)a.X;
    p.V = (const T *)a.V;
    p.W = (T *)a.W;
    p.H = (T *)a.H;
    p.hist = a.hist;
    p.niter = a.niter;
    p.n_rows = a.n_rows;
    p.ld = a.ld;
    p.m = a.m;
    p.max_iter = a.max_iter;
    p.update = a.update;
    {
        const char *e = getenv("CNMF_BATCHED_WPRE");
        p.wpre = e ? (e[0] != '0') : 1;
    }
    p.tol = a.tol;
    p.eps = (T)a.eps;
    kern<<<a.batch, kBT, smem, st>>>(p);
    count_launch();
    return (int)cudaGetLastError();
}

template <typename T, int K> static int launch_bk(const BatchedArgs &a, cudaStream_t st) {
    int bg, occ;
    batched_variant(K, sizeof(T) == 8 ? 1 : 0, &bg, &occ);
    if constexpr (K <= 4) {
        if (bg == 8 && occ == 2) return launch_bv<T, K, 8, 2>(a, st);
        if (bg == 8 && occ == 3) return launch_bv<T, K, 8, 3>(a, st);
        if (bg == 8 && occ == 4) return launch_bv<T, K, 8, 4>(a, st);
        if (bg == 4 && occ == 3) return launch_bv<T, K, 4, 3>(a, st);
        if (bg == 16 && occ == 3) return launch_bv<T, K, 16, 3>(a, st);
        if (bg == 16 && occ == 4) return launch_bv<T, K, 16, 4>(a, st);
    }
    return launch_bv<T, K, 4, 2>(a, st);
}

template <typename T> static int launch_bt(const BatchedArgs &a, cudaStream_t st) {
    switch (a.k) {
#define CASE(K) case K: return launch_bk<T, K>(a, st);
        CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7) CASE(8)
        CASE(9) CASE(10) CASE(11) CASE(12) CASE(13) CASE(14) CASE(15) CASE(16)
#undef CASE
    }
    return (int)cudaErrorInvalidValue;
}

int launch_batched(const BatchedArgs &a, cudaStream_t st) {
    return a.dtype == 0 ? launch_bt<float>(a, st) : launch_bt<double>(a, st);
}

}  // namespace cnmf
