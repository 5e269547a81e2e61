// K1, symmetric variant -- every worker warp does BOTH jobs of the fused pass.
//
// The role-split kernel (stream_kernel.cuh) leaves the FP pipe ~40 % idle: its five
// column-owner warps cannot be spread evenly over the four SM sub-partitions, they are the
// critical path, and the row-owner warps spend a fifth of their time waiting for stages.
// Here the NST x NWS worker warps are all alike:
//
//   * a worker OWNS a strip of columns (C2 packs per lane) for the tiles of its stream
//     (tile sequence numbers n = t mod NST): it keeps H for the strip and the k x C2
//     numerator / denominator partials in registers for the whole kernel (job H);
//   * whenever the next tile of its stream is not ready for job H, it takes a chunk of
//     rows from the CTA's shared queue and produces W' for them (job W: G lanes per row).
//
// Scheduling is work-conserving: no worker ever blocks while either kind of work exists,
// and all sub-partitions carry the same mix.  Completion of job W for a stage is tracked
// by monotonic counters in shared memory (wdone[s]); the only mbarriers left are full[s]
// (bulk-copy completion) and empty[s] (stage released by the NWS strip owners).
#pragma once
#include "stream_kernel.cuh"

namespace cnmf {

template <typename T, int K, int G, int C2, int RP1, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) fused_kernel(const StreamParams<T> p) {
    typedef Pk<T> O;
    typedef typename O::P P;
    constexpr int SZ = sizeof(T);
    constexpr int PW = O::W;
    constexpr int EPV = 16 / SZ;
    constexpr int KP = (K + 1) & ~1;
    constexpr int KS = h_stride_vecs(K);
    constexpr int RPW = 32 / G;
    constexpr int NV = next_pow2(K);
    constexpr bool HALVING = NV <= G;
    constexpr int CR = RPW * RP1;  // rows per job-W chunk
    extern __shared__ __align__(128) unsigned char smem[];

    if (p.state->done) return;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int R = p.R, stages = p.stages;
    const long long ld = p.ld;
    const int ldv = (int)(ld * SZ / 16);
    const int rowb = (int)(ld * SZ);
    const StreamSmem L = stream_smem(K, ld, SZ, R, stages);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + L.bar_off);
    uint64_t *empty = full + 2 * stages;  // (the slot between them is the role-split kernel's wready)
    unsigned int *chunk_ctr = reinterpret_cast<unsigned int *>(empty + stages);
    volatile int *issued = reinterpret_cast<volatile int *>(chunk_ctr + 1);
    unsigned int *wdone = chunk_ctr + 2;
    unsigned char *Hs = smem + L.h_off;
    const int cpt = R / CR;
    const int NST = p.S2, NWS = p.NW2;  // tile streams, strip-owner warps per stream

    if (tid == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], NWS);
            wdone[s] = 0u;
        }
        *chunk_ctr = 0u;
        *issued = 0;
        mbar_fence_init();
        if (blockIdx.x == 0) p.state->k1_count += 1;
    }
    {
        const int n = K * (int)ld;
        for (int i = tid; i < n; i += blockDim.x) {
            const int a = i / (int)ld, j = i - a * (int)ld;
            const int q = j / EPV, e = j - q * EPV;
            *reinterpret_cast<T *>(Hs + ((long long)q * KS + a) * 16 + e * SZ) =
                (j < p.m) ? p.H[(long long)a * p.m + j] : T(0);
        }
    }
    __syncthreads();

    const uint32_t full_a = smem_u32(full), empty_a = smem_u32(empty), wdone_a = smem_u32(wdone);
    unsigned char *const stage0 = smem + L.stage_off;
    const int sbytes = L.stage_bytes;
    const int last_tile = p.n_tiles - 1;
    const int last_rows = (int)(p.n_rows - (long long)last_tile * R);
    auto tile_rows = [&](int tile) { return tile == last_tile ? last_rows : R; };
    const int my_tiles = p.n_tiles > (int)blockIdx.x ? (p.n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

    if (warp == 0) {
        // ------------------------------------------------------------- producer
        if (lane == 0) {
            int s = 0, ph = 0, n = 0;
            for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
                mbar_wait_a(empty_a + s * 8, ph ^ 1);
                const int rows = tile_rows(tile);
                const uint32_t xb = (uint32_t)(rows * rowb);
                const uint32_t wb = (uint32_t)(rows * K * SZ);
                const bool wbulk = (wb & 15u) == 0;
                const uint32_t st = smem_u32(stage0) + s * sbytes;
                mbar_arrive_expect_tx_a(full_a + s * 8, 2 * xb + (wbulk ? wb : 0));
                const long long r0 = (long long)tile * R;
                bulk_g2s_a(st + L.x_off, p.X + r0 * ld, xb, full_a + s * 8);
                bulk_g2s_a(st + L.v_off, p.V + r0 * ld, xb, full_a + s * 8);
                if (wbulk) bulk_g2s_a(st + L.win_off, p.W + r0 * K, wb, full_a + s * 8);
                *issued = ++n;
                if (++s == stages) { s = 0; ph ^= 1; }
            }
        }
        return;
    }
    const int wk = warp - 1;
    if (wk >= NST * NWS) return;
    const int strm = wk / NWS;        // tile stream of this worker
    const int strip = wk - strm * NWS;  // its column strip

    // ------------------------------------------------------------- job H state
    const int ci = strip * 32 + lane;  // lane within the strip set
    const bool active = ci < p.TPS;
    const int col0 = ci * C2 * PW;  // first column of this lane
    P Hr[K][C2], nH[K][C2], dH[K][C2];
#pragma unroll
    for (int a = 0; a < K; ++a)
#pragma unroll
        for (int c = 0; c < C2; ++c) {
            const int col = col0 + c * PW;
            const int q = col / EPV, e = col - q * EPV;
            Hr[a][c] = active ? *reinterpret_cast<const P *>(Hs + ((long long)q * KS + a) * 16 + e * SZ) : O::zero();
            nH[a][c] = O::zero();
            dH[a][c] = O::zero();
        }
    double *blk = p.part + (long long)(blockIdx.x * NST + strm) * p.part_stride;
    bool first = true;
    auto flush = [&]() {
        if (active) {
#pragma unroll
            for (int a = 0; a < K; ++a)
#pragma unroll
                for (int c = 0; c < C2; ++c)
#pragma unroll
                    for (int e = 0; e < PW; ++e) {
                        const long long idx = (long long)a * ld + col0 + c * PW + e;
                        double nu = (double)O::lane(nH[a][c], e), de = (double)O::lane(dH[a][c], e);
                        if (!first) {
                            nu += blk[idx];
                            de += blk[(long long)K * ld + idx];
                        }
                        blk[idx] = nu;
                        blk[(long long)K * ld + idx] = de;
                    }
#pragma unroll
            for (int a = 0; a < K; ++a)
#pragma unroll
                for (int c = 0; c < C2; ++c) {
                    nH[a][c] = O::zero();
                    dH[a][c] = O::zero();
                }
        }
        first = false;
    };
    double chid = 0.0;

    // job H on the tile with sequence number n (all of its rows, this lane's columns)
    auto job_h = [&](int n) {
        const int pass = n / stages, s = n - pass * stages;
        const int tile = (int)blockIdx.x + n * (int)gridDim.x;
        const int rows = tile_rows(tile);
        const unsigned char *st = stage0 + s * sbytes;
        if (active && !(p.debug & 1)) {
            const unsigned char *xp = st + L.x_off + col0 * SZ;
            const unsigned char *vp = st + L.v_off + col0 * SZ;
            const unsigned char *wp = st + L.wd_off;
            P chi[C2];
#pragma unroll
            for (int c = 0; c < C2; ++c) chi[c] = O::zero();
            auto rows_u = [&](auto u_tag) {
                constexpr int UU = decltype(u_tag)::value;
                P x[UU][C2], v[UU][C2], wd[UU][KP];
#pragma unroll
                for (int u = 0; u < UU; ++u) {
                    if constexpr (C2 == 2) {
                        const Vec<T> xv = lds_vec<T>(xp + u * rowb);
                        const Vec<T> vv = lds_vec<T>(vp + u * rowb);
                        x[u][0] = xv.p[0]; x[u][1] = xv.p[1];
                        v[u][0] = vv.p[0]; v[u][1] = vv.p[1];
                    } else {
                        x[u][0] = *reinterpret_cast<const P *>(xp + u * rowb);
                        v[u][0] = *reinterpret_cast<const P *>(vp + u * rowb);
                    }
#pragma unroll
                    for (int a = 0; a < KP; a += 2) {
                        const Vec<T> t = lds_vec<T>(wp + u * (KP * 8) + a * 8);
                        wd[u][a] = t.p[0];
                        wd[u][a + 1] = t.p[1];
                    }
                }
                xp += UU * rowb;
                vp += UU * rowb;
                wp += UU * (KP * 8);
                P vx[UU][C2], wh[UU][C2];
#pragma unroll
                for (int u = 0; u < UU; ++u)
#pragma unroll
                    for (int c = 0; c < C2; ++c) {
                        vx[u][c] = O::mul(v[u][c], x[u][c]);
                        wh[u][c] = O::mul(wd[u][0], Hr[0][c]);
                    }
#pragma unroll
                for (int a = 1; a < K; ++a)
#pragma unroll
                    for (int u = 0; u < UU; ++u)
#pragma unroll
                        for (int c = 0; c < C2; ++c) wh[u][c] = O::fma(wd[u][a], Hr[a][c], wh[u][c]);
#pragma unroll
                for (int u = 0; u < UU; ++u)
#pragma unroll
                    for (int c = 0; c < C2; ++c) {
                        const P res = O::sub(x[u][c], wh[u][c]);
                        const P vr = O::mul(v[u][c], res);
                        chi[c] = O::fma(vr, res, chi[c]);
                        wh[u][c] = O::mul(v[u][c], wh[u][c]);  // V * W'H
                    }
#pragma unroll
                for (int u = 0; u < UU; ++u)
#pragma unroll
                    for (int a = 0; a < K; ++a)
#pragma unroll
                        for (int c = 0; c < C2; ++c) {
                            nH[a][c] = O::fma(wd[u][a], vx[u][c], nH[a][c]);
                            dH[a][c] = O::fma(wd[u][a], wh[u][c], dH[a][c]);
                        }
            };
            int i = 0;
#pragma unroll 1
            for (; i + 2 <= rows; i += 2) rows_u(std::integral_constant<int, 2>());
            if (i < rows) rows_u(std::integral_constant<int, 1>());
#pragma unroll
            for (int c = 0; c < C2; ++c) chid += O::dsum(chi[c]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive_a(empty_a + s * 8);
    };

    // ------------------------------------------------------------- job W on chunk c
    const int sub = lane % G;
    const int slot = HALVING ? row_reduce_slot<NV, G>(sub) : 0;
    const bool writer = HALVING ? ((sub & (G / NV - 1)) == 0 && slot < K) : false;
    auto job_w = [&](int n, int part) {
        const int pass = n / stages, s = n - pass * stages;
        const int tile = (int)blockIdx.x + n * (int)gridDim.x;
        const int rb = part * CR + lane / G;
        const int rows = tile_rows(tile);
        unsigned char *st = stage0 + s * sbytes;
        const bool wbulk = ((rows * K * SZ) & 15) == 0;
        bool valid[RP1];
        int rr[RP1];
        const T *win[RP1];
        P *wdst[RP1];
#pragma unroll
        for (int j = 0; j < RP1; ++j) {
            const int r = rb + j * RPW;
            valid[j] = r < rows;
            rr[j] = valid[j] ? r : 0;
            win[j] = wbulk ? reinterpret_cast<const T *>(st + L.win_off) + rr[j] * K
                           : p.W + ((long long)tile * R + rr[j]) * K;
            wdst[j] = reinterpret_cast<P *>(st + L.wd_off) + rr[j] * KP;
        }
        if (p.update_w) {
            P wd[RP1][K], num[RP1][K], den[RP1][K];
#pragma unroll
            for (int j = 0; j < RP1; ++j)
#pragma unroll
                for (int a = 0; a < K; ++a) {
                    wd[j][a] = O::dup(win[j][a]);
                    num[j][a] = O::zero();
                    den[j][a] = O::zero();
                }
            const unsigned char *xq[RP1], *vq[RP1];
#pragma unroll
            for (int j = 0; j < RP1; ++j) {
                xq[j] = st + L.x_off + rr[j] * rowb + sub * 16;
                vq[j] = st + L.v_off + rr[j] * rowb + sub * 16;
            }
            const unsigned char *hq = Hs + sub * (KS * 16);
            const int nq = (ldv - sub + G - 1) / G;
#pragma unroll 1
            for (int i = 0; i < nq; ++i) {
                Vec<T> h[K];
#pragma unroll
                for (int a = 0; a < K; ++a) h[a] = lds_vec<T>(hq + a * 16);
                hq += G * KS * 16;
                P vx[RP1][2], wh[RP1][2], vv[RP1][2];
#pragma unroll
                for (int j = 0; j < RP1; ++j) {
                    const Vec<T> x = lds_vec<T>(xq[j]);
                    const Vec<T> v = lds_vec<T>(vq[j]);
                    xq[j] += G * 16;
                    vq[j] += G * 16;
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        vv[j][e] = v.p[e];
                        vx[j][e] = O::mul(v.p[e], x.p[e]);
                        wh[j][e] = O::mul(wd[j][0], h[0].p[e]);
                    }
                }
#pragma unroll
                for (int a = 1; a < K; ++a)
#pragma unroll
                    for (int e = 0; e < 2; ++e)
#pragma unroll
                        for (int j = 0; j < RP1; ++j) wh[j][e] = O::fma(wd[j][a], h[a].p[e], wh[j][e]);
#pragma unroll
                for (int j = 0; j < RP1; ++j)
#pragma unroll
                    for (int e = 0; e < 2; ++e) wh[j][e] = O::mul(vv[j][e], wh[j][e]);  // V * WH
#pragma unroll
                for (int e = 0; e < 2; ++e)
#pragma unroll
                    for (int a = 0; a < K; ++a)
#pragma unroll
                        for (int j = 0; j < RP1; ++j) {
                            num[j][a] = O::fma(vx[j][e], h[a].p[e], num[j][a]);
                            den[j][a] = O::fma(wh[j][e], h[a].p[e], den[j][a]);
                        }
            }
#pragma unroll
            for (int j = 0; j < RP1; ++j) {
                const long long grow = (long long)tile * R + rr[j];
                if constexpr (HALVING) {
                    T nu[NV], de[NV];
#pragma unroll
                    for (int a = 0; a < NV; ++a) {
                        nu[a] = a < K ? (T)O::hsum(num[j][a < K ? a : 0]) : T(0);
                        de[a] = a < K ? (T)O::hsum(den[j][a < K ? a : 0]) : T(0);
                    }
                    RowReduce<T, NV, G>::run(nu, de, sub);
                    if (valid[j] && writer) {
                        const T wc = win[j][slot];
                        const T wn = (wc * nu[0]) / (de[0] > p.eps ? de[0] : p.eps);
                        p.W[grow * K + slot] = wn;
                        wdst[j][slot] = O::dup(wn);
                    }
                } else {
#pragma unroll
                    for (int a = 0; a < K; ++a) {
                        T nu = O::hsum(num[j][a]), de = O::hsum(den[j][a]);
#pragma unroll
                        for (int o = G / 2; o > 0; o >>= 1) {
                            nu += shfl_x(nu, o);
                            de += shfl_x(de, o);
                        }
                        if (valid[j] && (a % G) == sub) {
                            const T wn = (O::first(wd[j][a]) * nu) / (de > p.eps ? de : p.eps);
                            p.W[grow * K + a] = wn;
                            wdst[j][a] = O::dup(wn);
                        }
                    }
                }
            }
        } else {
#pragma unroll
            for (int j = 0; j < RP1; ++j)
                if (valid[j])
                    for (int a = sub; a < K; a += G) wdst[j][a] = O::dup(win[j][a]);
        }
        __syncwarp();
        if (lane == 0) red_release_add_a(wdone_a + s * 4);
    };

    // ------------------------------------------------------------- the worker's scheduler
    const unsigned total_chunks = (unsigned)my_tiles * (unsigned)cpt;
    constexpr unsigned NONE = 0xffffffffu;
    int hn = strm;          // next tile (sequence number) this worker owes job H
    unsigned c = NONE;      // claimed row chunk, not yet processed
    bool chunks_left = total_chunks > 0;
    int since = 0;
    while (true) {
        // 1. job H first: it releases stages and is what the producer waits for
        if (hn < my_tiles) {
            const int pass = hn / stages, s = hn - pass * stages;
            if (lds_acquire_u32_a(wdone_a + s * 4) >= (unsigned)(pass + 1) * (unsigned)cpt) {
                job_h(hn);
                hn += NST;
                if (SZ == 4 && ++since == p.flush_every) {
                    flush();
                    since = 0;
                }
                continue;
            }
        }
        // 2. otherwise a row chunk
        if (c == NONE && chunks_left) {
            unsigned g = 0;
            if (lane == 0) g = atomicAdd(chunk_ctr, 1u);
            g = __shfl_sync(0xffffffffu, g, 0);
            if (g < total_chunks) c = g;
            else chunks_left = false;
        }
        if (c != NONE) {
            const int n = (int)(c / (unsigned)cpt);
            // the tile must have been issued before full[s] may be waited on by parity (a wait
            // may be at most one phase ahead of its barrier)
            int is = 0;
            if (lane == 0) is = *issued;
            is = __shfl_sync(0xffffffffu, is, 0);
            if (is > n) {
                const int pass = n / stages, s = n - pass * stages;
                if (mbar_try_a(full_a + s * 8, pass & 1)) {
                    job_w(n, (int)(c - (unsigned)n * (unsigned)cpt));
                    c = NONE;
                }
            } else {
                __nanosleep(40);
            }
            continue;
        }
        if (hn >= my_tiles) break;  // no chunk left to claim, every owed tile done
        __nanosleep(40);            // only job H left and its tile is not complete yet
    }
    flush();
    chid = warp_sum(chid);
    if (lane == 0) p.chi_part[blockIdx.x * (NST * NWS) + wk] = chid;
}

}  // namespace cnmf
