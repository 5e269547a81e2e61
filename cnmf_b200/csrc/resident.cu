// K4r -- resident small-problem solver: one thread-block CLUSTER per independent problem, X and V
// loaded ONCE into the cluster's shared memory and kept there for all iterations.
//
// The streaming batched kernel (batched.cu) re-reads X and V from L2 every pass; a problem of
// config 4 (512 x 81) is 344 KB in fp32 / 672 KB in fp64, more than one SM's shared memory but not
// more than a cluster's: CS = 1, 2, 4 or 8 CTAs (the smallest that fits) each keep n_rows / CS rows
// of X and V, their W rows (as the duplicated packs phase 2 reads) and a full copy of H.  Per pass a
// CTA runs phase 1 (G lanes per row -> W') and phase 2 (thread owns one pack of columns -> chi2 and
// the k x M partials) over its resident rows, sums what its column streams accumulated in fp64 (fixed
// order), and the CTAs of the cluster then read each other's sums through distributed shared memory
// (in rank order, so every CTA applies the bit-identical H update to its own copy).  The sums are
// double-buffered, so ONE cluster barrier and three CTA barriers make an iteration; there is no
// global-memory traffic inside the loop at all.
// The same kernel is the small single-problem path (config 1: 1 000 x 81 fp64 = 8 CTAs x 125 rows).
#include <cooperative_groups.h>
#include <stdio.h>
#include <stdlib.h>

#include "launch.h"
#include "phases.cuh"

namespace cg = cooperative_groups;

namespace cnmf {

constexpr int kRT = 512;  // threads per CTA: one CTA per SM (shared memory), 16 warps to hide latency

template <typename T> struct ResidentParams {
    const T *X;
    const T *V;
    T *W;
    T *H;
    double *hist;
    int *niter;
    long long n_rows, ld;
    int m, max_iter, update, rows_per;
    double tol;
    T eps;
};

struct ResidentSmem {
    int h_off, acc_off, acc_bytes, red_off, sl_off, wd_off, x_off, v_off, total;
};
__host__ __device__ inline ResidentSmem resident_smem(int K, long long ld, int sz, int rows_per) {
    ResidentSmem s;
    auto up = [](long long v) { return (int)((v + 127) / 128 * 128); };
    const int KP = (K + 1) & ~1;
    const int packs = (int)(ld / (sz == 4 ? 2 : 1));
    const int streams = packs > 0 && packs <= kRT ? kRT / packs : 1;
    s.h_off = 0;                              // [K][ld] T
    s.acc_off = up((long long)K * ld * sz);   // [2 buffers][2][K][ld] fp64 sums over this CTA's rows (peers read them)
    s.acc_bytes = up(2LL * K * ld * 8);
    s.red_off = s.acc_off + 2 * s.acc_bytes;  // [16] warp chi2 sums, [16], [17] this CTA's chi2 (two buffers)
    s.sl_off = s.red_off + 256;               // [streams][2][K][ld] T: what each column stream accumulated
    s.wd_off = s.sl_off + up((long long)streams * 2 * K * ld * sz);  // [rows_per][KP] W rows as duplicated packs
    s.x_off = s.wd_off + up((long long)rows_per * KP * 8);
    s.v_off = s.x_off + up((long long)rows_per * ld * sz);
    s.total = s.v_off + up((long long)rows_per * ld * sz);
    return s;
}

__device__ __forceinline__ void r_cp_async16(void *dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}

template <typename T, int K, int BG>
__global__ void __launch_bounds__(kRT, 1) resident_kernel(const ResidentParams<T> p) {
    typedef Pk<T> O;
    typedef typename O::P P;
    constexpr int SZ = sizeof(T);
    constexpr int PW = O::W;
    constexpr int KP = (K + 1) & ~1;
    constexpr int RPS = kRT / BG;  // rows per phase-1 sweep
    extern __shared__ __align__(128) unsigned char smem[];
    cg::cluster_group cluster = cg::this_cluster();
    const int cs = (int)cluster.num_blocks(), rank = (int)cluster.block_rank();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long ld = p.ld;
    const int ldi = (int)ld;
    const int ldv = (int)(ld * SZ / 16);
    const int packs = (int)(ld / PW);
    const ResidentSmem L = resident_smem(K, ld, SZ, p.rows_per);
    T *Hs = reinterpret_cast<T *>(smem + L.h_off);
    double *red = reinterpret_cast<double *>(smem + L.red_off);
    T *slices = reinterpret_cast<T *>(smem + L.sl_off);
    P *Wd = reinterpret_cast<P *>(smem + L.wd_off);
    unsigned char *Xs = smem + L.x_off, *Vs = smem + L.v_off;
    const long long b = blockIdx.x / cs;
    const long long my_r0 = (long long)rank * p.rows_per;
    const int my_rows = (int)(p.n_rows - my_r0 < 0 ? 0 : (p.n_rows - my_r0 < p.rows_per ? p.n_rows - my_r0 : p.rows_per));
    const T *X = p.X + (b * p.n_rows + my_r0) * ld;
    const T *V = p.V + (b * p.n_rows + my_r0) * ld;
    T *W = p.W + (b * p.n_rows + my_r0) * K;
    T *Hg = p.H + b * (long long)K * p.m;
    double *hist = p.hist + b * (long long)(p.max_iter + 1);
    const int streams = kRT / packs;  // >= 1 (checked on the host)
    const int stream = tid / packs, ci = tid - stream * packs;
    const bool active = stream < streams;
    const int r1 = tid / BG, sub = tid % BG;

    // ---- one-time loads: X, V (cp.async), W (as duplicated packs), H
    {
        const int nvec = my_rows * ldv;
        const unsigned char *xs = reinterpret_cast<const unsigned char *>(X);
        const unsigned char *vs = reinterpret_cast<const unsigned char *>(V);
        for (int i = tid; i < nvec; i += kRT) {
            r_cp_async16(Xs + i * 16, xs + (long long)i * 16);
            r_cp_async16(Vs + i * 16, vs + (long long)i * 16);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        for (int i = tid; i < my_rows * K; i += kRT) {
            const int r = i / K, a = i - r * K;
            Wd[r * KP + a] = O::dup(W[i]);
        }
        for (int i = tid; i < K * ldi; i += kRT) {
            const int a = i / ldi, j = i - a * ldi;
            Hs[i] = (j < p.m) ? Hg[(long long)a * p.m + j] : T(0);
        }
        if (rank == 0)
            for (int i = tid; i <= p.max_iter; i += kRT) hist[i] = __longlong_as_double(0x7ff8000000000000LL);
        asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();

#ifdef CNMF_PROF
    long long tph[6] = {0, 0, 0, 0, 0, 0}, tq = clock64();
#define RTICK(i) { const long long now_ = clock64(); tph[i] += now_ - tq; tq = now_; }
#else
#define RTICK(i)
#endif
    double prev_chi = 0.0;
    int it_done = 0;
    // pass 0 evaluates (W0, H0); pass it >= 1 updates W with the H the previous pass produced
    for (int pass = 0; pass <= p.max_iter; ++pass) {
        const bool upd_w = pass > 0 && (p.update & 1);
        if (upd_w) {  // phase 1 over the resident rows: NR rows per lane group, NR * RPS rows per sweep
            constexpr int NR = K <= 4 ? 2 : 1;  // (registers: 128 per thread)
            for (int rb = 0; rb < my_rows; rb += NR * RPS) {
                bool valid[NR];
                int rr[NR];
                const unsigned char *xr[NR], *vr[NR];
                T w[NR][K], wn[NR][K];
#pragma unroll
                for (int j = 0; j < NR; ++j) {
                    const int r = rb + j * RPS + r1;
                    valid[j] = r < my_rows;
                    rr[j] = valid[j] ? r : 0;
                    xr[j] = Xs + (long long)rr[j] * ld * SZ;
                    vr[j] = Vs + (long long)rr[j] * ld * SZ;
#pragma unroll
                    for (int a = 0; a < K; ++a) w[j][a] = O::first(Wd[rr[j] * KP + a]);
                }
                phase1_rows<T, K, BG, NR>(xr, vr, reinterpret_cast<const unsigned char *>(Hs), ldv, sub, w, p.eps, wn);
                __syncwarp();  // every lane of the rows read their W before one of them replaces it
#pragma unroll
                for (int j = 0; j < NR; ++j)
                    if (valid[j]) {
#pragma unroll
                        for (int a = 0; a < K; ++a)
                            if ((a % BG) == sub) Wd[rr[j] * KP + a] = O::dup(wn[j][a]);
                    }
            }
            __syncthreads();
        }
        RTICK(0)
        // phase 2: this thread's pack of columns over the rows of its stream
        double chid = 0.0;
        if (active) {
            P Hr[K], nH[K], dH[K];
#pragma unroll
            for (int a = 0; a < K; ++a) {
                Hr[a] = *reinterpret_cast<const P *>(reinterpret_cast<const unsigned char *>(Hs) + ((long long)a * ld + ci * PW) * SZ);
                nH[a] = O::zero();
                dH[a] = O::zero();
            }
            const unsigned char *xp = Xs + ((long long)stream * ld + ci * PW) * SZ;
            const unsigned char *vp = Vs + ((long long)stream * ld + ci * PW) * SZ;
            const long long step = (long long)streams * ld * SZ;
            // NR2 rows in flight: their loads are issued together, their W'H chains advance side by side
            constexpr int NR2 = K <= 4 ? 4 : 2;
            P chi[NR2];
#pragma unroll
            for (int u = 0; u < NR2; ++u) chi[u] = O::zero();
            int since = 0;
            int r = stream;
            for (; r + (NR2 - 1) * streams < my_rows; r += NR2 * streams) {
                P x[NR2], v[NR2], wd[NR2][K];
#pragma unroll
                for (int u = 0; u < NR2; ++u) {
                    x[u] = *reinterpret_cast<const P *>(xp + u * step);
                    v[u] = *reinterpret_cast<const P *>(vp + u * step);
#pragma unroll
                    for (int a = 0; a < K; ++a) wd[u][a] = Wd[(r + u * streams) * KP + a];
                }
                xp += NR2 * step;
                vp += NR2 * step;
#pragma unroll
                for (int u = 0; u < NR2; ++u) phase2_pack<T, K>(x[u], v[u], wd[u], Hr, nH, dH, chi[u]);
                if (SZ == 4 && (since += NR2) >= 16) {  // fp32: chi2 leaves the packed accumulators every 16 rows
#pragma unroll
                    for (int u = 0; u < NR2; ++u) {
                        chid += O::dsum(chi[u]);
                        chi[u] = O::zero();
                    }
                    since = 0;
                }
            }
            for (; r < my_rows; r += streams) {
                const P x0 = *reinterpret_cast<const P *>(xp), v0 = *reinterpret_cast<const P *>(vp);
                xp += step;
                vp += step;
                P wd0[K];
#pragma unroll
                for (int a = 0; a < K; ++a) wd0[a] = Wd[r * KP + a];
                phase2_pack<T, K>(x0, v0, wd0, Hr, nH, dH, chi[0]);
            }
#pragma unroll
            for (int u = 0; u < NR2; ++u) chid += O::dsum(chi[u]);
            // what this stream accumulated goes to its slice (summed in fp64, in stream order, below)
            P *sl = reinterpret_cast<P *>(slices + (long long)stream * 2 * K * ldi) + ci;
#pragma unroll
            for (int a = 0; a < K; ++a) {
                sl[a * packs] = nH[a];
                sl[(K + a) * packs] = dH[a];
            }
        }
        chid = warp_sum(chid);
        if (lane == 0) red[warp] = chid;
        __syncthreads();
        RTICK(1)
        // ---- this CTA's sums in fp64 (fixed order), into the buffer of this pass's parity
        double *acc = reinterpret_cast<double *>(smem + L.acc_off + (pass & 1) * L.acc_bytes);
        for (int i = tid; i < 2 * K * ldi; i += kRT) {
            double v = 0.0;
            for (int s2 = 0; s2 < streams; ++s2) v += (double)slices[(long long)s2 * 2 * K * ldi + i];
            acc[i] = v;
        }
        if (tid == 0) {
            double c = 0.0;
            for (int w = 0; w < kRT / 32; ++w) c += red[w];
            red[16 + (pass & 1)] = c;
        }
        // ---- the cluster's sums: every CTA reads every CTA's partials (rank order) and keeps its own H.
        // One cluster barrier per pass is enough: the sums are double-buffered, and a CTA reaches the
        // barrier of pass n + 1 only after it has read everybody's sums of pass n.
        RTICK(2)
        cluster.sync();
        RTICK(3)
        double c = 0.0;
        for (int q = 0; q < cs; ++q) c += cluster.map_shared_rank(red, q)[16 + (pass & 1)];
        if (rank == 0 && tid == 0) hist[pass] = c;
        bool stop = !(c == c) || c > 1.7e308;
        if (p.tol > 0.0 && pass >= 1 && (prev_chi - c) <= p.tol * prev_chi) stop = true;
        prev_chi = c;
        it_done = pass;
        if (stop || pass == p.max_iter) break;
        if (p.update & 2) {
            for (int i = tid; i < K * p.m; i += kRT) {
                const int a = i / p.m, j = i - a * p.m;
                double num = 0.0, den = 0.0;
                for (int q = 0; q < cs; ++q) {
                    const double *ra = cluster.map_shared_rank(acc, q);
                    num += ra[a * ldi + j];
                    den += ra[K * ldi + a * ldi + j];
                }
                const T nu = (T)num, de = (T)den;
                const T h = Hs[a * ldi + j];
                Hs[a * ldi + j] = (h * nu) / (de > p.eps ? de : p.eps);
            }
        }
        __syncthreads();  // H of the next pass is in place
        RTICK(4)
    }
#ifdef CNMF_PROF
    if (blockIdx.x == 0 && tid == 0)
        printf("resident cs=%d rows=%d passes=%d: cycles/pass phase1 %lld phase2 %lld combine %lld cluster-barrier %lld sums+H %lld\n", cs,
               my_rows, it_done + 1, tph[0] / (it_done + 1), tph[1] / (it_done + 1), tph[2] / (it_done + 1), tph[3] / (it_done + 1),
               tph[4] / (it_done + 1));
#endif
    cluster.sync();  // no CTA exits while a peer may still read its sums
    if (p.update & 1)
        for (int i = tid; i < my_rows * K; i += kRT) {
            const int r = i / K, a = i - r * K;
            W[i] = O::first(Wd[r * KP + a]);
        }
    if (rank == 0) {
        for (int i = tid; i < K * p.m; i += kRT) {
            const int a = i / p.m, j = i - a * p.m;
            Hg[i] = Hs[a * ldi + j];
        }
        if (tid == 0) p.niter[b] = it_done;
    }
}

// cluster size for a problem shape: the smallest of 1, 2, 4, 8 whose share of X and V fits one SM's
// shared memory; 0 if none does (or the columns do not fit one pack per thread)
int resident_cluster_size(int dtype, int k, long long ld, long long n_rows, int max_smem, int *rows_per_out,
                          int *smem_out) {
    const int sz = dtype == 0 ? 4 : 8, pw = dtype == 0 ? 2 : 1;
    if (k > 8 || ld / pw > kRT || n_rows < 1) return 0;
    for (int cs = 1; cs <= 8; cs *= 2) {
        const long long rows_per = (n_rows + cs - 1) / cs;
        if (rows_per * ld * sz * 2 > max_smem) continue;
        const ResidentSmem L = resident_smem(k, ld, sz, (int)rows_per);
        if (L.total <= max_smem) {
            if (rows_per_out) *rows_per_out = (int)rows_per;
            if (smem_out) *smem_out = L.total;
            return cs;
        }
    }
    return 0;
}

template <typename T, int K, int BG> static int launch_rb(const BatchedArgs &a, int cs, int rows_per, cudaStream_t st) {
    auto kern = resident_kernel<T, K, BG>;
    const int smem = resident_smem(K, a.ld, sizeof(T), rows_per).total;
    static int smem_set[kMaxDevices] = {0};  // the opt-in above 48 KB is a per-device function attribute
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= kMaxDevices) return (int)cudaErrorInvalidDevice;
    if (smem > smem_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return (int)e;
        smem_set[dev] = smem;
    }
    if (a.max_iter < 0) {
        // probe (resident_can_launch): can the device hold at least one cluster of this shape right now?
        cudaLaunchConfig_t q = {};
        q.gridDim = dim3((unsigned)cs, 1, 1);
        q.blockDim = dim3(kRT, 1, 1);
        q.dynamicSmemBytes = (size_t)smem;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = (unsigned)cs;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        q.attrs = at;
        q.numAttrs = 1;
        int n = 0;
        cudaError_t e = cudaOccupancyMaxActiveClusters(&n, kern, &q);
        if (e != cudaSuccess) {
            (void)cudaGetLastError();
            return -1;
        }
        return n >= 1 ? 0 : -1;
    }
    ResidentParams<T> p;
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
    p.rows_per = rows_per;
    p.tol = a.tol;
    p.eps = (T)a.eps;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(a.batch * cs), 1, 1);
    cfg.blockDim = dim3(kRT, 1, 1);
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)cs;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, p);
    count_launch();
    return (int)e;
}

template <typename T, int K> static int launch_rk(const BatchedArgs &a, int cs, int rows_per, cudaStream_t st) {
    if constexpr (sizeof(T) == 4) {
        // fp32: 4 lanes per row.  (Two rows then share a quarter-warp and their X / V vectors collide in
        // the banks at most row pitches; 8 lanes per row avoid that but measured slower -- config 4:
        // 0.469 vs 0.411 ms per iteration -- the 21 vectors of a row leave a quarter of them idle.)
        const char *e = getenv("CNMF_RES_BG");  // sweeps only
        if (e && atoi(e) == 8) return launch_rb<T, K, 8>(a, cs, rows_per, st);
        return launch_rb<T, K, 4>(a, cs, rows_per, st);
    }
    return launch_rb<T, K, 8>(a, cs, rows_per, st);
}

template <typename T> static int launch_rt(const BatchedArgs &a, int cs, int rows_per, cudaStream_t st) {
    switch (a.k) {
#define CASE(K) case K: return launch_rk<T, K>(a, cs, rows_per, st);
        CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7) CASE(8)
#undef CASE
    }
    return (int)cudaErrorInvalidValue;
}

// true if a cluster of `cs` CTAs of this problem shape can be co-scheduled on the current device
bool resident_can_launch(int dtype, int k, long long ld, int m, int cs, int rows_per) {
    BatchedArgs a = {};
    a.ld = ld;
    a.m = m;
    a.k = k;
    a.dtype = dtype;
    a.batch = 1;
    a.max_iter = -1;  // probe only
    return (a.dtype == 0 ? launch_rt<float>(a, cs, rows_per, nullptr) : launch_rt<double>(a, cs, rows_per, nullptr)) == 0;
}

int launch_resident(const BatchedArgs &a, int cs, int rows_per, cudaStream_t st) {
    return a.dtype == 0 ? launch_rt<float>(a, cs, rows_per, st) : launch_rt<double>(a, cs, rows_per, st);
}

}  // namespace cnmf
