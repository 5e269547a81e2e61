// K1, "eight fat warps" form of the fused row-streaming kernel (fp64, 256 < ld <= 384, k <= 6).
//
// The 12-warp form (stream_kernel.cuh) is register-starved: 168 registers per thread leave
// the compiler no room to keep the next rows' operands in flight, so every pass of a
// column-owner warp starts with an exposed shared-memory round trip, and five column-owner
// warps on four SM sub-partitions leave one sub-partition with two of them.  Here a CTA is
// 8 warps x 255 registers:
//   warps 0-3  column owners, ONE per sub-partition.  Two or three columns per lane (n3 warps
//              own 96 columns, the others 64, so four warps cover up to 384), row pairs
//              software-pipelined through two register sets: the loads of pair i + 1 are issued
//              before the arithmetic of pair i.
//   warps 4-7  row owners (8 lanes per row, 4-row chunks from the shared queue), as in the
//              12-warp kernel.
//   no producer warp: the column-owner warp that releases a stage LAST (shared counter) arms
//              the stage's barrier and issues the bulk copies of the tile that reuses it.
// Data layout, partial blocks and barriers are those of stream_kernel.cuh.
#pragma once
#include "stream_kernel.cuh"

namespace cnmf {

__device__ __forceinline__ double lds_f64(const void *p) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(smem_u32(p)));
    return v;
}

constexpr int kFatThreads = 256;

template <int K>
__global__ void __launch_bounds__(kFatThreads, 1) fat_kernel(const StreamParams<double> p) {
    typedef double T;
    constexpr int SZ = 8;
    constexpr int KP = (K + 1) & ~1;
    constexpr int KS = h_stride_vecs(K);
    constexpr int G = 8, RPW = 4;  // lanes per row, rows per row-owner chunk
    constexpr int NV = next_pow2(K);
    constexpr int NWH = 4;
    static_assert(NV <= G, "recursive halving needs k <= 8");
    extern __shared__ __align__(128) unsigned char smem[];

    if (p.state->done) return;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int R = p.R, stages = p.stages;
    const long long ld = p.ld;
    const int ldv = (int)(ld / 2);
    const int rowb = (int)(ld * SZ);
    const StreamSmem L = stream_smem(K, ld, SZ, R, stages);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + L.bar_off);
    uint64_t *wready = full + stages;
    // the 12-warp kernel's empty[] barriers become two counters per stage
    unsigned int *armed = reinterpret_cast<unsigned int *>(wready + stages);  // times stage s was armed and its copies issued
    unsigned int *freed = armed + stages;                                      // column-owner arrivals on stage s
    unsigned int *chunk_ctr = freed + stages;
    unsigned int *wdone = chunk_ctr + 2;
    unsigned char *Hs = smem + L.h_off;
    const int cpt = R / RPW;
    const int kFatZeroOff = fat_zero_off(stages);  // idle lanes load X = V = 0 from here

    if (tid == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&wready[s], cpt);
            armed[s] = 0u;
            freed[s] = 0u;
            wdone[s] = 0u;
        }
        *chunk_ctr = 0u;
        mbar_fence_init();
        if (blockIdx.x == 0) p.state->k1_count += 1;
    }
    if (tid < 4) reinterpret_cast<unsigned int *>(smem + kFatZeroOff)[tid] = 0u;
    {
        const int n = K * (int)ld;
        for (int i = tid; i < n; i += kFatThreads) {
            const int a = i / (int)ld, j = i - a * (int)ld;
            const int q = j >> 1, e = j & 1;
            *reinterpret_cast<T *>(Hs + ((long long)q * KS + a) * 16 + e * SZ) = (j < p.m) ? p.H[(long long)a * p.m + j] : T(0);
        }
    }
    __syncthreads();

    const uint32_t full_a = smem_u32(full), wready_a = smem_u32(wready);
    const uint32_t wdone_a = smem_u32(wdone), freed_a = smem_u32(freed);
    volatile unsigned int *armed_v = armed;
    unsigned char *const stage0 = smem + L.stage_off;
    const int sbytes = L.stage_bytes;
    const int last_tile = p.n_tiles - 1;
    const int last_rows = (int)(p.n_rows - (long long)last_tile * R);
    auto tile_rows = [&](int tile) { return tile == last_tile ? last_rows : R; };
    const int my_tiles = p.n_tiles > (int)blockIdx.x ? (p.n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

    // arm stage n % stages for the n-th tile of this CTA and issue its copies (one lane)
    auto issue = [&](int n) {
        const int pass = n / stages, s = n - pass * stages;
        const int tile = (int)blockIdx.x + n * (int)gridDim.x;
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
        armed_v[s] = (unsigned)pass + 1u;
    };
    if (tid == 0) {
        const int n0 = my_tiles < stages ? my_tiles : stages;
        for (int n = 0; n < n0; ++n) issue(n);
    }

    if (warp >= NWH) {
        // ------------------------------------------------- row owners (as stream_kernel, G = 8, one row per lane group)
        const int sub = lane % G;
        const int slot = row_reduce_slot<NV, G>(sub);
        const bool writer = ((sub & (G / NV - 1)) == 0 && slot < K);
        const unsigned total_chunks = (unsigned)my_tiles * (unsigned)cpt;
        auto grab = [&]() {
            unsigned c = 0;
            if (lane == 0) c = atomicAdd(chunk_ctr, 1u);
            return __shfl_sync(0xffffffffu, c, 0);
        };
        unsigned c = grab();
        while (c < total_chunks) {
            const unsigned cnext = grab();
            const int n = (int)(c / (unsigned)cpt);
            const int part = (int)(c - (unsigned)n * (unsigned)cpt);
            const int pass = n / stages;
            const int s = n - pass * stages;
            const int tile = (int)blockIdx.x + n * (int)gridDim.x;
            // chunks are claimed ahead of the loads: wait until the stage was armed for this pass,
            // then the parity wait cannot alias with an older phase
            while (!__all_sync(0xffffffffu, armed_v[s] > (unsigned)pass)) __nanosleep(32);
            mbar_wait_warp_a(full_a + s * 8, pass & 1);
            {
                const int r = part * RPW + lane / G;
                const int rows = tile_rows(tile);
                unsigned char *st = stage0 + s * sbytes;
                const bool wbulk = ((rows * K * SZ) & 15) == 0;
                const bool valid = r < rows;
                const int rr = valid ? r : 0;
                const T *win = wbulk ? reinterpret_cast<const T *>(st + L.win_off) + rr * K : p.W + ((long long)tile * R + rr) * K;
                T *wdst = reinterpret_cast<T *>(st + L.wd_off) + rr * KP;
                if (p.update_w) {
                    T wd[K], num[K], den[K];
#pragma unroll
                    for (int a = 0; a < K; ++a) {
                        wd[a] = win[a];
                        num[a] = 0.0;
                        den[a] = 0.0;
                    }
                    const unsigned char *xq = st + L.x_off + rr * rowb + sub * 16;
                    const unsigned char *vq = st + L.v_off + rr * rowb + sub * 16;
                    const unsigned char *hq = Hs + sub * (KS * 16);
                    const int nq = (ldv - sub + G - 1) / G;
                    auto vecs = [&](auto vu_tag) {
                        constexpr int VU = decltype(vu_tag)::value;
                        Vec<T> h[VU][K];
                        T vx[VU][2], wh[VU][2], vv[VU][2];
#pragma unroll
                        for (int u = 0; u < VU; ++u) {
#pragma unroll
                            for (int a = 0; a < K; ++a) h[u][a] = lds_vec<T>(hq + u * (G * KS * 16) + a * 16);
                            const Vec<T> x = lds_vec<T>(xq + u * (G * 16));
                            const Vec<T> v = lds_vec<T>(vq + u * (G * 16));
#pragma unroll
                            for (int e = 0; e < 2; ++e) {
                                vv[u][e] = v.p[e];
                                vx[u][e] = v.p[e] * x.p[e];
                                wh[u][e] = wd[0] * h[u][0].p[e];
                            }
                        }
                        hq += VU * (G * KS * 16);
                        xq += VU * (G * 16);
                        vq += VU * (G * 16);
#pragma unroll
                        for (int a = 1; a < K; ++a)
#pragma unroll
                            for (int u = 0; u < VU; ++u)
#pragma unroll
                                for (int e = 0; e < 2; ++e) wh[u][e] = fma(wd[a], h[u][a].p[e], wh[u][e]);
#pragma unroll
                        for (int u = 0; u < VU; ++u)
#pragma unroll
                            for (int e = 0; e < 2; ++e) wh[u][e] = vv[u][e] * wh[u][e];
#pragma unroll
                        for (int u = 0; u < VU; ++u)
#pragma unroll
                            for (int e = 0; e < 2; ++e)
#pragma unroll
                                for (int a = 0; a < K; ++a) {
                                    num[a] = fma(vx[u][e], h[u][a].p[e], num[a]);
                                    den[a] = fma(wh[u][e], h[u][a].p[e], den[a]);
                                }
                    };
                    int i = 0;
#pragma unroll 1
                    for (; i + 2 <= nq; i += 2) vecs(std::integral_constant<int, 2>());
                    if (i < nq) vecs(std::integral_constant<int, 1>());
                    T nu[NV], de[NV];
#pragma unroll
                    for (int a = 0; a < NV; ++a) {
                        nu[a] = a < K ? num[a < K ? a : 0] : T(0);
                        de[a] = a < K ? den[a < K ? a : 0] : T(0);
                    }
                    RowReduce<T, NV, G>::run(nu, de, sub);
                    if (valid && writer) {
                        const T wc = win[slot];
                        const T wn = (wc * nu[0]) / (de[0] > p.eps ? de[0] : p.eps);
                        p.W[((long long)tile * R + rr) * K + slot] = wn;
                        wdst[slot] = wn;
                    }
                } else if (valid) {
                    for (int a = sub; a < K; a += G) wdst[a] = win[a];
                }
                __syncwarp();
                if (lane == 0) {
                    red_release_add_a(wdone_a + s * 4);
                    mbar_arrive_a(wready_a + s * 8);
                }
            }
            c = cnext;
        }
    } else {
        // ---------------------------------------------- column owners: one warp per sub-partition
        const int hw = warp;
        const int n3 = ((int)ld - 256 + 31) / 32;  // warps that own 96 columns (host guarantees 0 <= n3 <= 4)
        const bool three = hw < n3;
        const int base = three ? hw * 96 : n3 * 96 + (hw - n3) * 64;
        const int colA = base + 2 * lane;  // columns colA, colA + 1
        const int colB = base + 64 + lane;  // third column of a 96-column warp
        const bool actA = colA < (int)ld;
        const bool actB = three && colB < (int)ld;
        double *blk = p.part + (long long)blockIdx.x * p.part_stride;
        double chid = 0.0;

        auto run = [&](auto nc_tag) {
            constexpr int NC = decltype(nc_tag)::value;
            T Hr[K][NC], nH[K][NC], dH[K][NC];
#pragma unroll
            for (int a = 0; a < K; ++a) {
                const Vec<T> t = lds_vec<T>(Hs + ((long long)((actA ? colA : 0) >> 1) * KS + a) * 16);
                Hr[a][0] = actA ? t.p[0] : 0.0;
                Hr[a][1] = actA ? t.p[1] : 0.0;
                if constexpr (NC == 3) {
                    const int cb = actB ? colB : 0;
                    Hr[a][2] = actB ? *reinterpret_cast<const T *>(Hs + ((long long)(cb >> 1) * KS + a) * 16 + (cb & 1) * 8) : 0.0;
                }
#pragma unroll
                for (int c = 0; c < NC; ++c) {
                    nH[a][c] = 0.0;
                    dH[a][c] = 0.0;
                }
            }
            T chi[NC];
#pragma unroll
            for (int c = 0; c < NC; ++c) chi[c] = 0.0;
            // idle lanes read zeros with a zero row step
            const int stepA = actA ? rowb : 0, stepB = actB ? rowb : 0;
            // X and V of a row pair live in one of two register sets (the loads of pair i + 1 are
            // issued before the arithmetic of pair i); W' has one set, reloaded row by row right
            // behind the accumulations that read it last
            struct Set {
                T x[2][NC], v[2][NC];
            };
            T wd[2][KP];
            int s = 0, ph = 0, pass = 0;
            bool ready = false;
            for (int n = 0; n < my_tiles; ++n) {
                const int tile = (int)blockIdx.x + n * (int)gridDim.x;
                if (!ready) mbar_wait_warp_a(wready_a + s * 8, ph);
                const int s_next = (s + 1 == stages) ? 0 : s + 1;
                const int pass_next = (s + 1 == stages) ? pass + 1 : pass;
                const unsigned done_next = lds_acquire_u32_a(wdone_a + s_next * 4);
                const int rows = tile_rows(tile);
                const unsigned char *st = stage0 + s * sbytes;
                if (!(p.debug & 1)) {
                    const unsigned char *xa = actA ? st + L.x_off + colA * SZ : smem + kFatZeroOff;
                    const unsigned char *va = actA ? st + L.v_off + colA * SZ : smem + kFatZeroOff;
                    const unsigned char *xb = actB ? st + L.x_off + colB * SZ : smem + kFatZeroOff;
                    const unsigned char *vb = actB ? st + L.v_off + colB * SZ : smem + kFatZeroOff;
                    const unsigned char *wp = st + L.wd_off;
                    auto load = [&](Set &t) {
#pragma unroll
                        for (int u = 0; u < 2; ++u) {
                            const Vec<T> xv = lds_vec<T>(xa + u * stepA);
                            const Vec<T> vv = lds_vec<T>(va + u * stepA);
                            t.x[u][0] = xv.p[0];
                            t.x[u][1] = xv.p[1];
                            t.v[u][0] = vv.p[0];
                            t.v[u][1] = vv.p[1];
                            if constexpr (NC == 3) {
                                t.x[u][2] = lds_f64(xb + u * stepB);
                                t.v[u][2] = lds_f64(vb + u * stepB);
                            }
                        }
                        xa += 2 * stepA;
                        va += 2 * stepA;
                        xb += 2 * stepB;
                        vb += 2 * stepB;
                    };
                    auto load_wd = [&](int u) {
#pragma unroll
                        for (int a = 0; a < KP; a += 2) {
                            const Vec<T> w = lds_vec<T>(wp + u * (KP * 8) + a * 8);
                            wd[u][a] = w.p[0];
                            wd[u][a + 1] = w.p[1];
                        }
                    };
                    // `more`: another pair of this tile follows (its W' is fetched here)
                    auto compute = [&](const Set &t, auto uu_tag, bool more) {
                        constexpr int UU = decltype(uu_tag)::value;
                        T vx[UU][NC], wh[UU][NC];
#pragma unroll
                        for (int u = 0; u < UU; ++u)
#pragma unroll
                            for (int c = 0; c < NC; ++c) {
                                vx[u][c] = t.v[u][c] * t.x[u][c];
                                wh[u][c] = wd[u][0] * Hr[0][c];
                            }
#pragma unroll
                        for (int a = 1; a < K; ++a)
#pragma unroll
                            for (int u = 0; u < UU; ++u)
#pragma unroll
                                for (int c = 0; c < NC; ++c) wh[u][c] = fma(wd[u][a], Hr[a][c], wh[u][c]);
#pragma unroll
                        for (int u = 0; u < UU; ++u)
#pragma unroll
                            for (int c = 0; c < NC; ++c) {
                                const T res = t.x[u][c] - wh[u][c];
                                const T vr = t.v[u][c] * res;
                                chi[c] = fma(vr, res, chi[c]);
                                wh[u][c] = t.v[u][c] * wh[u][c];
                            }
                        wp += 2 * (KP * 8);
#pragma unroll
                        for (int u = 0; u < UU; ++u) {
#pragma unroll
                            for (int a = 0; a < K; ++a)
#pragma unroll
                                for (int c = 0; c < NC; ++c) {
                                    nH[a][c] = fma(wd[u][a], vx[u][c], nH[a][c]);
                                    dH[a][c] = fma(wd[u][a], wh[u][c], dH[a][c]);
                                }
                            if (more) load_wd(u);
                        }
                    };
                    const int pairs = rows >> 1;
                    Set A, B;
                    load_wd(0);
                    load_wd(1);
                    if (pairs > 0) load(A);
                    int it = 0;
#pragma unroll 1
                    for (; it + 2 <= pairs; it += 2) {
                        load(B);
                        compute(A, std::integral_constant<int, 2>(), true);
                        const bool more = it + 2 < pairs;
                        if (more) load(A);
                        compute(B, std::integral_constant<int, 2>(), more || (rows & 1));
                    }
                    if (pairs & 1) compute(A, std::integral_constant<int, 2>(), (rows & 1) != 0);
                    if (rows & 1) {
                        // last row of a ragged last tile: the pointers stand on it (row u = 1 of
                        // these loads lies past the tile, inside the stage; it is not used)
                        load(A);
                        compute(A, std::integral_constant<int, 1>(), false);
                    }
                }
                __syncwarp();
                if (lane == 0) {
                    const unsigned old = atom_add_acqrel_a(freed_a + s * 4);
                    if (((old + 1u) & (NWH - 1)) == 0u && n + stages < my_tiles) issue(n + stages);
                }
                ready = done_next >= (unsigned)(pass_next + 1) * (unsigned)cpt;
                s = s_next;
                if (s == 0) ph ^= 1;
                pass = pass_next;
            }
            // this CTA's partial block: num[k][ld] then den[k][ld]
#pragma unroll
            for (int a = 0; a < K; ++a) {
                if (actA) {
                    blk[(long long)a * ld + colA] = nH[a][0];
                    blk[(long long)a * ld + colA + 1] = nH[a][1];
                    blk[(long long)(K + a) * ld + colA] = dH[a][0];
                    blk[(long long)(K + a) * ld + colA + 1] = dH[a][1];
                }
                if constexpr (NC == 3) {
                    if (actB) {
                        blk[(long long)a * ld + colB] = nH[a][2];
                        blk[(long long)(K + a) * ld + colB] = dH[a][2];
                    }
                }
            }
#pragma unroll
            for (int c = 0; c < NC; ++c) chid += chi[c];
        };
        if (three) run(std::integral_constant<int, 3>());
        else run(std::integral_constant<int, 2>());
        chid = warp_sum(chid);
        if (lane == 0) p.chi_part[blockIdx.x * NWH + hw] = chid;
    }
}

}  // namespace cnmf
