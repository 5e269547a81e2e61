// K1 -- the fused row-streaming kernel (>95 % of the time of an iteration).
//
// One persistent CTA per SM walks row tiles of X and V.  Each tile is brought
// into shared memory ONCE by the TMA engine (cp.async.bulk, one flat copy per
// operand: rows are contiguous because ld is the pitch) through an mbarrier
// ring, and is consumed by two warp-specialised roles that overlap on
// different tiles:
//
//   role W (phase 1, "row owners"): G lanes share a row; with H in shared
//       memory they form WH on the fly and accumulate (V*X)H^T and (V*WH)H^T
//       in registers, reduce the 2k sums over the G lanes by recursive halving
//       (lane c ends up owning component c), apply
//       W' = W * num / max(den, eps), store W' to HBM and (as duplicated
//       packs) to shared memory for role H.
//   role H (phase 2, "column owners"): each thread owns C2 packs of columns,
//       keeps H for them and the k x C2 numerator / denominator partials
//       W'^T(V*X), W'^T(V*W'H) in registers for the whole kernel, walks the
//       rows of the tile, and also accumulates chi2 = sum V (X - W'H)^2.
//
// Per element that is 6k+7 FMA-pipe operations; fp32 issues
// them two columns at a time as FFMA2.  The k x M partials leave the kernel
// as fp64 blocks (one per CTA and H-stream); fp32 accumulators are flushed
// to them every `flush_every` tiles so a 10^7-row sum keeps fp64 accuracy.
#pragma once
#include <stdio.h>
#include <type_traits>

#include "common.cuh"

namespace cnmf {

// -DCNMF_PROF builds (cnmf_b200/build.py --prof -> libcnmf_b200_prof.so) time the phases of
// each role with clock64 and print them for CTA 0; the product library carries none of it.
#ifdef CNMF_PROF
#define PROF_DECL(...) long long __VA_ARGS__
#define PROF_T(v) const long long v = clock64()
#define PROF_ADD(acc, t0) acc += clock64() - (t0)
// resource probes of the diagnostics build (CNMF_TUNE DBG bits 2..4, results invalid): which of the
// shared-memory loads / FMAs a role can lose before the kernel time moves
#define PROBE_OFF(bit) (!(p.debug & (bit)))
#define PROBE_ZERO = {}
#else
#define PROF_DECL(...)
#define PROF_T(v)
#define PROF_ADD(acc, t0)
#define PROBE_OFF(bit) true
#define PROBE_ZERO
#endif

template <typename T> struct StreamParams {
    const T *X;
    const T *V;
    T *W;
    const T *H;
    double *part;      // [grid * S2][part_stride]: num[k][ld] then den[k][ld]
    double *chi_part;  // [grid * NW2]
    State *state;
    long long n_rows;
    long long ld;
    long long part_stride;
    int m;
    int n_tiles;
    int R;       // rows per tile
    int NW1;     // row-owner warps (take row chunks of the tiles from a shared counter)
    int NW2;     // column-owner warps
    int S2;      // column-owner streams (rows of a tile are dealt round-robin)
    int TPS;     // threads per stream
    int stages;
    int update_w;
    int flush_every;
    int debug;  // diagnostics build only (-DCNMF_PROF): probe bits, see PROBE_OFF; the product kernels never read it
    int prefetch;  // L2 prefetch distance of the producer, in tiles of this CTA (0 = off)
    int NT;        // the last (partly filled) column-owner warp exists NT times, each taking every NT-th of its rows
    int crows;     // NI > 0 only: rows of a role-W chunk (taken one after the other), <= 32 / next_pow2(K)
    int noprod;    // 1: no producer warp -- the column-owner warp that frees a stage last issues the next copies into it;
                   // 2: no producer warp -- lane 31 (idle) of the last column-owner warp runs the producer loop on its own
    int gate_ns;          // tuning builds: sleep of the armed[] gate poll (product: 32 ns)
    int lookahead;        // tuning builds: 0 = role H always waits on wready[] (no wdone[] look-ahead)
    unsigned stages_inv;  // ceil(2^32 / stages): tile index -> (ring pass, stage) without an integer division
    T eps;
    // role of each warp: 0 = producer, 1 + i = row owner i, 128 + i = column-owner warp i.
    // The host deals roles to warps so that the FMA work of the four SM sub-partitions
    // (warp % 4) is balanced (api.cu, deal_roles).
    unsigned char role[32];
    // mma_kernel.cuh: first column group and number of column groups of each column-owner warp
    unsigned char hg0[16], hgn[16];
};

// H in shared memory is stored vector-major: the 16-byte vector q of row a sits
// at (q * KS + a) * 16 with KS = K rounded up to odd, so that the G lanes of a
// row (consecutive q) hit distinct banks and role W addresses H[a] with an
// immediate offset from one moving pointer.
__host__ __device__ constexpr int h_stride_vecs(int K) { return K | 1; }

// offset of 16 zero bytes behind the barrier words (idle lanes of fat_kernel load X = V = 0 there)
__host__ __device__ constexpr int fat_zero_off(int stages) { return (3 * stages * 8 + 8 + 8 * stages + 15) / 16 * 16; }

// shared-memory carve-up, shared by host (sizing) and device
struct StreamSmem {
    int bar_off, h_off, stage_off, stage_bytes, x_off, v_off, win_off, wd_off, total;
};
__host__ __device__ inline StreamSmem stream_smem(int K, long long ld, int sz, int R, int stages) {
    StreamSmem s;
    auto up = [](long long v, long long a) { return (int)((v + a - 1) / a * a); };
    const int KP = (K + 1) & ~1;
    const long long ldv = ld * sz / 16;
    s.bar_off = 0;
    // barriers, chunk counter, issued count, per-stage chunk-done counts, 16 zero bytes (fat_kernel.cuh)
    s.h_off = up(fat_zero_off(stages) + 16, 128);
    s.stage_off = up(s.h_off + ldv * h_stride_vecs(K) * 16, 128);
    s.x_off = 0;
    s.v_off = up((long long)R * ld * sz, 128);
    s.win_off = s.v_off + up((long long)R * ld * sz, 128);
    s.wd_off = s.win_off + up((long long)R * K * sz, 128);
    s.stage_bytes = s.wd_off + up((long long)R * KP * 8, 128);
    s.total = s.stage_off + stages * s.stage_bytes;
    return s;
}

// 64-bit-safe select / shuffle helpers for the recursive-halving reduction
template <typename T> __device__ __forceinline__ T shfl_x(T v, int mask) {
    return __shfl_xor_sync(0xffffffffu, v, mask);
}

// Reduce NV (power of two) slot pairs (num, den) over the G lanes of a row so
// that the lane with sub-index s holds the totals of slot (s * NV / G) when
// NV >= G ... in general slot = s >> log2(G / NV') after halving stops.
// Implemented for the two shapes used: values are halved while more than one
// slot remains, the remaining lane bits are summed with a butterfly.
template <typename T, int NV, int G> struct RowReduce {
    // in: n[NV], d[NV]; out: n[0], d[0] = totals of slot `slot_of(sub)`
    static __device__ __forceinline__ void run(T (&n)[NV], T (&d)[NV], int sub) {
        constexpr int HALF = NV / 2;
        if constexpr (NV > 1 && G > 1) {
            const bool hi = (sub & (G / 2)) != 0;
            T n2[HALF], d2[HALF];
#pragma unroll
            for (int i = 0; i < HALF; ++i) {
                const T sn = hi ? n[i] : n[i + HALF];
                const T sd = hi ? d[i] : d[i + HALF];
                const T kn = hi ? n[i + HALF] : n[i];
                const T kd = hi ? d[i + HALF] : d[i];
                n2[i] = kn + shfl_x(sn, G / 2);
                d2[i] = kd + shfl_x(sd, G / 2);
            }
            RowReduce<T, HALF, G / 2>::run(n2, d2, sub);
            n[0] = n2[0];
            d[0] = d2[0];
        } else if constexpr (G > 1) {
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) {
                n[0] += shfl_x(n[0], o);
                d[0] += shfl_x(d[0], o);
            }
        }
    }
};
// slot owned by lane `sub` after RowReduce<T, NV, G>
template <int NV, int G> __device__ __forceinline__ int row_reduce_slot(int sub) {
    // each halving step consumes the current top lane bit and picks the upper half of slots if set
    int slot = 0, nv = NV, g = G;
    while (nv > 1 && g > 1) {
        nv >>= 1;
        if (sub & (g >> 1)) slot += nv;
        g >>= 1;
    }
    return slot;
}

// W' = W num / max(den, eps): den >= eps > 0 and finite.  fp64: reciprocal seed + two Newton steps + one
// correction of the quotient (within 1 ulp of the IEEE quotient; the inlined IEEE division with its slow
// path is ~60 instructions per chunk of role W).  fp32: the hardware sequence of the compiler.
__device__ __forceinline__ double ratio_div(double num, double den) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(den));
    r = fma(fma(-den, r, 1.0), r, r);
    r = fma(fma(-den, r, 1.0), r, r);
    const double q = num * r;
    return fma(fma(-den, q, num), r, q);
}
__device__ __forceinline__ float ratio_div(float num, float den) { return num / den; }

__host__ __device__ constexpr int next_pow2(int k) { return k <= 1 ? 1 : (k <= 2 ? 2 : (k <= 4 ? 4 : (k <= 8 ? 8 : 16))); }

// NI > 0 selects the register-resident form of role W (G = 32, RP1 = 1): a row-owner warp takes ONE
// row at a time, lane l owns the packs l, l + 32, ... (NI = ceil(packs / 32) of them) and keeps H
// for them in registers for the whole kernel, so role W reads only X and V from shared memory
// (0.125 instead of 0.5 wavefronts per fp64 element); the 2k sums cross the 32 lanes by recursive
// halving and the W' arithmetic of the p.crows rows of a chunk is done once, one (row, component)
// per lane.
template <typename T, int K, int G, int C2, int RP1, int MAXT, int NI = 0>
__global__ void __launch_bounds__(MAXT, 1) stream_kernel(const StreamParams<T> p) {
    typedef Pk<T> O;
    typedef typename O::P P;
    constexpr int SZ = sizeof(T);
    constexpr int PW = O::W;
    constexpr int EPV = 16 / SZ;  // elements per 16-byte vector
    constexpr int KP = (K + 1) & ~1;
    constexpr int KS = h_stride_vecs(K);
    constexpr int RPW = 32 / G;  // rows per row-owner warp
    constexpr int NV = next_pow2(K);
    constexpr bool HALVING = NV <= G;  // recursive-halving reduction leaves one component per lane
    extern __shared__ __align__(128) unsigned char smem[];

    if (p.state->done) return;  // a previous finalize hit the tolerance: become a no-op

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int R = p.R, stages = p.stages;
#ifdef CNMF_TUNABLE
    const int noprod = p.noprod;  // producer-less variants exist in the tuning builds only
#else
    constexpr int noprod = 0;
#endif
    const long long ld = p.ld;
    const int ldv = (int)(ld * SZ / 16);  // 16-byte vectors per row
    const int rowb = (int)(ld * SZ);      // bytes per row
    const StreamSmem L = stream_smem(K, ld, SZ, R, stages);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + L.bar_off);
    uint64_t *wready = full + stages;
    uint64_t *empty = wready + stages;
    unsigned int *chunk_ctr = reinterpret_cast<unsigned int *>(empty + stages);
    // armed[s]: how many times stage s was armed and its copies issued (row owners claim chunks
    // ahead of the loads and gate on it).  Without a producer warp the empty[] barriers are not
    // used and their words hold armed[] and freed[] (column-owner arrivals per stage); with one,
    // armed[] lives behind wdone[].
    volatile unsigned int *armed = noprod == 1 ? reinterpret_cast<volatile unsigned int *>(empty)
                                            : reinterpret_cast<volatile unsigned int *>(chunk_ctr + 2 + stages);
    unsigned int *freed = reinterpret_cast<unsigned int *>(empty) + stages;
    // wdone[s]: row chunks finished in stage s since the kernel started.  Role H reads it with
    // a plain (asynchronous) shared load one tile ahead, so that in the common case -- role W
    // is ahead -- it never pays the ~100-cycle round trip of an mbarrier wait.
    unsigned int *wdone = chunk_ctr + 2;
    unsigned char *Hs = smem + L.h_off;
    static_assert(NI == 0 || (G == 32 && RP1 == 1), "register-resident role W: one row per warp pass");
    const int CR = NI > 0 ? p.crows : RPW * RP1;  // rows per row-owner chunk
    const int cpt = R / CR;                       // chunks per tile
    const int role = p.role[warp];

    if (tid == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&wready[s], cpt);
            if (noprod != 1) mbar_init(&empty[s], p.NW2);
        }
        *chunk_ctr = 0u;
        for (int s = 0; s < stages; ++s) {
            wdone[s] = 0u;
            armed[s] = 0u;
            if (noprod == 1) freed[s] = 0u;
        }
        mbar_fence_init();
        if (blockIdx.x == 0) p.state->k1_count += 1;
    }
    // H (k x m, dense) -> shared (vector-major, see h_stride_vecs), zero in the pad columns
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

    const uint32_t full_a = smem_u32(full), wready_a = smem_u32(wready), empty_a = smem_u32(empty);
    const uint32_t wdone_a = smem_u32(wdone), freed_a = smem_u32(freed);
    unsigned char *const stage0 = smem + L.stage_off;
    const int sbytes = L.stage_bytes;
    const int last_tile = p.n_tiles - 1;
    const int last_rows = (int)(p.n_rows - (long long)last_tile * R);
    auto tile_rows = [&](int tile) { return tile == last_tile ? last_rows : R; };
    // tiles this CTA walks: blockIdx.x, blockIdx.x + gridDim.x, ...
    const int my_tiles = p.n_tiles > (int)blockIdx.x ? (p.n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

    // arm stage n % stages for the n-th tile of this CTA and issue its bulk copies (one lane)
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
        armed[s] = (unsigned)pass + 1u;
    };
    if (noprod == 1 && tid == 0) {
        const int n0 = my_tiles < stages ? my_tiles : stages;
        for (int n = 0; n < n0; ++n) issue(n);
    }

    // the producer loop (one lane): refill each stage as soon as the column owners released it
    auto producer_loop = [&]() {
        int s = 0, ph = 0, n = 0;
        PROF_DECL(pw = 0);
        PROF_T(pt0);
        const int pf = p.prefetch;  // tiles of L2 prefetch distance (0 = off)
        for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
            if (pf > 0) {
                const long long tpf = (long long)tile + (long long)pf * gridDim.x;
                if (tpf < p.n_tiles) {
                    const int rows = tile_rows((int)tpf);
                    bulk_prefetch_l2(p.X + tpf * R * ld, (uint32_t)(rows * rowb));
                    bulk_prefetch_l2(p.V + tpf * R * ld, (uint32_t)(rows * rowb));
                }
            }
            PROF_T(pa);
            mbar_wait_a(empty_a + s * 8, ph ^ 1);
            PROF_ADD(pw, pa);
            issue(n);
            ++n;
            if (++s == stages) { s = 0; ph ^= 1; }
        }
#ifdef CNMF_PROF
        if (blockIdx.x == 0) printf("warp %2d producer: total %lld wait-empty %lld tiles %d\n", warp, clock64() - pt0, pw, n);
#endif
    };

    if (role == 0) {
        // ------------------------------------------------------------- producer
        if (lane == 0) producer_loop();
    } else if (role < 128) {
        // ------------------------------------------------- role W: row owners
        // The row-owner warps share one queue of row chunks (CR rows of a tile, in tile
        // order): a warp on a lightly loaded sub-partition simply takes more of them.
        // A lane works on RP1 rows at once (rows rb, rb + RPW, ...): every H vector it
        // loads from shared memory is used for all of them.
        const int sub = lane % G;
        // component this lane owns after the reduction, and whether it is the writer
        const int slot = HALVING ? row_reduce_slot<NV, G>(sub) : 0;
        const bool writer = HALVING ? ((sub & (G / NV - 1)) == 0 && slot < K) : false;
        const unsigned total_chunks = (unsigned)my_tiles * (unsigned)cpt;
        const int cpt_sh = 31 - __clz(cpt);
        // register-resident form: H of this lane's packs, and the byte offsets of its packs in a row
        // (the last pack index is clamped for the lanes past the end of the row; their H is zero)
        constexpr int NIA = NI > 0 ? NI : 1;
        P Hw[NIA][K];
        int wr_lane8 = lane * 8, wr_last = 0;
        if constexpr (NI > 0) {
            const int packs = (int)ld / PW;
#pragma unroll
            for (int i = 0; i < NI; ++i) {
                const int pk = lane + 32 * i;
                const bool act = pk < packs;
                const int col = (act ? pk : 0) * PW;
                const int q = col / EPV, e = col - q * EPV;
                if (i + 1 == NI) wr_last = (act ? pk : packs - 1) * 8 - wr_lane8;  // relative to the lane's first pack
#pragma unroll
                for (int a = 0; a < K; ++a)
                    Hw[i][a] = act ? *reinterpret_cast<const P *>(Hs + ((long long)q * KS + a) * 16 + e * SZ) : O::zero();
            }
        }
        auto grab = [&]() {
            unsigned c = 0;
            if (lane == 0) c = atomicAdd(chunk_ctr, 1u);
            return __shfl_sync(0xffffffffu, c, 0);
        };
        PROF_DECL(wt_gate = 0, wt_full = 0, wt_loop = 0, wt_tail = 0, wn = 0);
        PROF_T(wt0);
        unsigned c = grab();
        while (c < total_chunks) {
            const unsigned cnext = grab();  // claimed one chunk ahead: hides the atomic's latency
            PROF_T(wa);
            // cpt is a power of two (the plan sees to it) and stages_inv = ceil(2^32 / stages): no integer division
            const int n = (int)(c >> cpt_sh);
            const int part = (int)(c & (unsigned)(cpt - 1));
            const int pass = (int)__umulhi((unsigned)n, p.stages_inv);
            const int s = n - pass * stages;
            const int tile = (int)blockIdx.x + n * (int)gridDim.x;
            // A parity wait may only be one phase ahead of its barrier, but chunks are claimed
            // ahead of the loads (by up to two per warp, and bulk copies may land out of
            // order).  Gate on the producer's count of issued tiles first: once tile n was
            // issued, full[s] is in phase `pass` (pending or complete) and the parity wait
            // below cannot alias with an older phase.
            // (every lane polls and the exit is a vote: a lane-0-only spin left lane 0 running the
            // whole chunk as a warp of its own, see mbar_wait_warp_a)
#ifdef CNMF_PROF
            long long wb = wa;
#endif
#ifdef CNMF_TUNABLE
            while (!__all_sync(0xffffffffu, armed[s] > (unsigned)pass)) __nanosleep(p.gate_ns);
#else
            // (a pass that does not update W leaves the row owners with a copy per chunk: they poll at a
            // longer interval so that their polling does not take issue slots from the column owners)
            while (!__all_sync(0xffffffffu, armed[s] > (unsigned)pass)) __nanosleep(p.update_w ? 32 : 400);
#endif
            PROF_ADD(wt_gate, wa);
#ifdef CNMF_PROF
            wb = clock64();
#endif
            mbar_wait_warp_a(full_a + s * 8, pass & 1);
            PROF_ADD(wt_full, wb);
            PROF_T(wc);
#ifdef CNMF_PROF
            long long wc2 = wc;
            ++wn;
#endif
            if constexpr (NI > 0) {
                const int rows = tile_rows(tile);
                unsigned char *st = stage0 + s * sbytes;
                const bool wbulk = ((rows * K * SZ) & 15) == 0;
                constexpr int LPS = 32 / NV;  // lanes that end up holding the totals of one component
                const int r0 = part * CR;
                auto w_in = [&](int r) {
                    return wbulk ? reinterpret_cast<const T *>(st + L.win_off) + r * K : p.W + ((long long)tile * R + r) * K;
                };
                if (p.update_w) {
                    T myn = T(0), myd = T(1);
#pragma unroll 1
                    for (int j = 0; j < CR; ++j) {
                        const int r = r0 + j < rows ? r0 + j : 0;  // rows past the end redo row 0, nothing is stored
                        const T *win = w_in(r);
                        P wd[K], num[K], den[K];
#pragma unroll
                        for (int a = 0; a < K; ++a) {
                            wd[a] = O::dup(win[a]);
                            num[a] = O::zero();
                            den[a] = O::zero();
                        }
                        // one moving pointer per operand, the packs at immediate offsets from it
                        const unsigned char *xq = st + L.x_off + r * rowb + wr_lane8;
                        const unsigned char *vq = st + L.v_off + r * rowb + wr_lane8;
#pragma unroll
                        for (int i = 0; i < NI; ++i) {
                            const P x = *reinterpret_cast<const P *>(i + 1 < NI ? xq + i * 256 : xq + wr_last);
                            const P v = *reinterpret_cast<const P *>(i + 1 < NI ? vq + i * 256 : vq + wr_last);
                            const P vx = O::mul(v, x);
                            P wh = O::mul(wd[0], Hw[i][0]);
#pragma unroll
                            for (int a = 1; a < K; ++a) wh = O::fma(wd[a], Hw[i][a], wh);
                            wh = O::mul(v, wh);  // V * WH
#pragma unroll
                            for (int a = 0; a < K; ++a) {
                                num[a] = O::fma(vx, Hw[i][a], num[a]);
                                den[a] = O::fma(wh, Hw[i][a], den[a]);
                            }
                        }
                        T nu[NV], de[NV];
#pragma unroll
                        for (int a = 0; a < NV; ++a) {
                            nu[a] = a < K ? (T)O::hsum(num[a < K ? a : 0]) : T(0);
                            de[a] = a < K ? (T)O::hsum(den[a < K ? a : 0]) : T(0);
                        }
                        RowReduce<T, NV, 32>::run(nu, de, lane);
                        if ((lane & (LPS - 1)) == j) {  // this lane finishes row j of the chunk
                            myn = nu[0];
                            myd = de[0];
                        }
                    }
#ifdef CNMF_PROF
                    wc2 = clock64();
                    wt_loop += wc2 - wc;
#endif
                    const int t = lane & (LPS - 1);
                    if (t < CR && r0 + t < rows && slot < K) {
                        const int r = r0 + t;
                        const T wcur = w_in(r)[slot];
                        const T wn = (wcur * myn) / (myd > p.eps ? myd : p.eps);
                        p.W[((long long)tile * R + r) * K + slot] = wn;
                        (reinterpret_cast<P *>(st + L.wd_off) + r * KP)[slot] = O::dup(wn);
                    }
                } else {
                    for (int i = lane; i < CR * K; i += 32) {
                        const int j = i / K, a = i - j * K;
                        if (r0 + j < rows)
                            (reinterpret_cast<P *>(st + L.wd_off) + (r0 + j) * KP)[a] = O::dup(w_in(r0 + j)[a]);
                    }
                }
                __syncwarp();
                if (lane == 0) {
                    red_release_add_a(wdone_a + s * 4);
                    mbar_arrive_a(wready_a + s * 8);
                }
            } else {
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
                    // VU vectors of a row per pass: VU x RP1 x 2 chains WH advance together (a chain
                    // step is a dependent FMA of 8 (fp64) / 4 (fp32) cycles; four chains cover it)
                    auto vecs = [&](auto vu_tag) {
                        constexpr int VU = decltype(vu_tag)::value;
                        Vec<T> h[VU][K] PROBE_ZERO;
                        P vx[VU][RP1][2], wh[VU][RP1][2], vv[VU][RP1][2];
#pragma unroll
                        for (int u = 0; u < VU; ++u) {
#pragma unroll
                            for (int a = 0; a < K; ++a)
                                if (PROBE_OFF(4)) h[u][a] = lds_vec<T>(hq + u * (G * KS * 16) + a * 16);  // bit 2: probe without the H loads
#pragma unroll
                            for (int j = 0; j < RP1; ++j) {
                                const Vec<T> x = lds_vec<T>(xq[j] + u * (G * 16));
                                const Vec<T> v = lds_vec<T>(vq[j] + u * (G * 16));
#pragma unroll
                                for (int e = 0; e < 2; ++e) {
                                    vv[u][j][e] = v.p[e];
                                    vx[u][j][e] = O::mul(v.p[e], x.p[e]);
                                    wh[u][j][e] = O::mul(wd[j][0], h[u][0].p[e]);
                                }
                            }
                        }
                        hq += VU * (G * KS * 16);
#pragma unroll
                        for (int j = 0; j < RP1; ++j) {
                            xq[j] += VU * (G * 16);
                            vq[j] += VU * (G * 16);
                        }
#pragma unroll
                        for (int a = 1; a < K; ++a)
#pragma unroll
                            for (int u = 0; u < VU; ++u)
#pragma unroll
                                for (int e = 0; e < 2; ++e)
#pragma unroll
                                    for (int j = 0; j < RP1; ++j)
                                        wh[u][j][e] = O::fma(wd[j][a], h[u][a].p[e], wh[u][j][e]);
#pragma unroll
                        for (int u = 0; u < VU; ++u)
#pragma unroll
                            for (int j = 0; j < RP1; ++j)
#pragma unroll
                                for (int e = 0; e < 2; ++e) wh[u][j][e] = O::mul(vv[u][j][e], wh[u][j][e]);  // V * WH
                        if (PROBE_OFF(8))  // bit 3: probe without the 2k accumulation FMAs
#pragma unroll
                        for (int u = 0; u < VU; ++u)
#pragma unroll
                            for (int e = 0; e < 2; ++e)
#pragma unroll
                                for (int a = 0; a < K; ++a)
#pragma unroll
                                    for (int j = 0; j < RP1; ++j) {
                                        num[j][a] = O::fma(vx[u][j][e], h[u][a].p[e], num[j][a]);
                                        den[j][a] = O::fma(wh[u][j][e], h[u][a].p[e], den[j][a]);
                                    }
                    };
                    {
                        constexpr int VUMAX = RP1 == 1 ? 2 : 1;
                        int i = 0;
                        if constexpr (VUMAX == 2) {
#pragma unroll 1
                            for (; i + 2 <= nq; i += 2) vecs(std::integral_constant<int, 2>());
                        }
#pragma unroll 1
                        for (; i < nq; ++i) vecs(std::integral_constant<int, 1>());
                    }
#ifdef CNMF_PROF
                    wc2 = clock64();
                    wt_loop += wc2 - wc;
#endif
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
                                const T wn = ratio_div(wc * nu[0], de[0] > p.eps ? de[0] : p.eps);
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
                                    const T wn = ratio_div(O::first(wd[j][a]) * nu, de > p.eps ? de : p.eps);
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
                if (lane == 0) {
                    red_release_add_a(wdone_a + s * 4);
                    mbar_arrive_a(wready_a + s * 8);
                }
            }
            PROF_ADD(wt_tail, wc2);
            c = cnext;
        }
#ifdef CNMF_PROF
        if (blockIdx.x == 0 && lane == 0)
            printf("warp %2d role W: total %lld gate %lld full %lld loop %lld tail %lld chunks %lld\n", warp, clock64() - wt0,
                   wt_gate, wt_full, wt_loop, wt_tail, wn);
#endif
    } else if (role - 128 < p.NW2) {
        // ---------------------------------------------- role H: column owners
        const int w2 = role - 128;
        // The S2 * TPS column-owner lanes rarely fill whole warps (M = 301: 151 lanes = 4.72 warps),
        // and a warp with 23 live lanes costs its sub-partition as much as a full one.  The last
        // warp therefore exists NT times, each copy walking every NT-th of its rows: its cost is
        // spread over NT sub-partitions (warp indices nfull .. NW2-1 all map to lane block nfull).
        const int nfull = p.NW2 - p.NT;
        const bool tail = w2 >= nfull;
        const int tp = tail ? w2 - nfull : 0;
        const int t2 = (tail ? nfull : w2) * 32 + lane;
        const int stream = t2 / p.TPS;
        const int ci = t2 - stream * p.TPS;
        const bool active = stream < p.S2;
        const int pc = ci * C2;    // first pack of this thread
        const int col0 = pc * PW;  // first column
        const int S2 = tail ? p.S2 * p.NT : p.S2;           // row step of this lane
        const int rphase = stream + (tail ? p.S2 * tp : 0);  // its first row within a tile
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
        // partial block: one per row phase; blocks of tail phases > 0 only ever see the tail's
        // columns (the plan zero-fills the scratch once)
        double *blk = p.part + (long long)(blockIdx.x * (p.S2 * p.NT) + (active ? rphase : 0)) * p.part_stride;
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
                            // Every entry of a block belongs to ONE lane: the first flush of a launch overwrites what the last launch
                            // left, later ones add with a fire-and-forget reduction (RED.ADD.F64; same thread, same address: the adds
                            // keep their order, so the sum is bit-reproducible).  As a load-add-store chain every entry cost an L2
                            // round trip -- the compiler may not hoist a load over an earlier store to the same array -- which was
                            // 5 % (k = 6, 301 columns) to 7 % (k = 8, 184 columns) of an fp32 iteration (profiles/r2c_flush_interval.txt).
                            const double nu = (double)O::lane(nH[a][c], e), de = (double)O::lane(dH[a][c], e);
                            if (first) {
                                blk[idx] = nu;
                                blk[(long long)K * ld + idx] = de;
                            } else {
                                atomicAdd(blk + idx, nu);
                                atomicAdd(blk + (long long)K * ld + idx, de);
                            }
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
        const int xstep = S2 * rowb, wstep = S2 * KP * 8;
        const int nr_full = active && rphase < R ? (R - rphase + S2 - 1) / S2 : 0;
        // byte offsets of this thread's first row inside a stage
        const int xoff0 = L.x_off + col0 * SZ + rphase * rowb, voff0 = L.v_off + col0 * SZ + rphase * rowb;
        const int woff0 = L.wd_off + rphase * (KP * 8);
        int s = 0, ph = 0, pass = 0, since = 0, nloc = 0;
        bool ready = false;  // the coming tile was already seen complete through wdone[]
        PROF_DECL(ht_wait = 0, ht_loop = 0, hn = 0, hn_ready = 0);
        PROF_T(ht0);
        // noprod == 2: lane 31 of the last column-owner warp (an idle lane: the host checks it) runs
        // the producer loop as a diverged path of its own; the warp's collectives leave it out.
        const bool carrier = noprod == 2 && w2 == p.NW2 - 1;
        const unsigned hmask = carrier ? 0x7fffffffu : 0xffffffffu;
        if (carrier && lane == 31) {
            producer_loop();
        } else {
        for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
            // wready[s] completes after role W saw full[s]: the tile is visible through that chain
            PROF_T(ha);
#ifdef CNMF_PROF
            ++hn;
            hn_ready += ready;
#endif
            if (!ready) mbar_wait_warp_a(wready_a + s * 8, ph, hmask);
            PROF_ADD(ht_wait, ha);
            PROF_T(hb);
            // look one tile ahead: the count is consumed after this tile's rows
            const int s_next = (s + 1 == stages) ? 0 : s + 1;
            const int pass_next = (s + 1 == stages) ? pass + 1 : pass;
            const unsigned done_next = lds_acquire_u32_a(wdone_a + s_next * 4);
            // only the last tile of the problem can be short: its row count is worked out off the common path
            int nr = nr_full;
            if (tile == last_tile && last_rows != R) nr = active && rphase < last_rows ? (last_rows - rphase + S2 - 1) / S2 : 0;
            const unsigned char *st = stage0 + s * sbytes;
            if (nr > 0 && PROBE_OFF(1)) {  // bit 0 (diagnostics build): role H skips its arithmetic
                const unsigned char *xp = st + xoff0;
                const unsigned char *vp = st + voff0;
                const unsigned char *wp = st + woff0;
                P chi[C2];
#pragma unroll
                for (int c = 0; c < C2; ++c) chi[c] = O::zero();
                // UU rows at a time: their UU x C2 chains W'H advance together
                auto rows_u = [&](auto u_tag) {
                    constexpr int UU = decltype(u_tag)::value;
                    P x[UU][C2], v[UU][C2], wd[UU][KP];
#pragma unroll
                    for (int u = 0; u < UU; ++u) {
                        if constexpr (C2 == 2) {
                            Vec<T> xv PROBE_ZERO, vv PROBE_ZERO;
                            if (PROBE_OFF(16)) {  // bit 4: probe without role H's X / V loads
                                xv = lds_vec<T>(xp + u * xstep);
                                vv = lds_vec<T>(vp + u * xstep);
                            }
                            x[u][0] = xv.p[0]; x[u][1] = xv.p[1];
                            v[u][0] = vv.p[0]; v[u][1] = vv.p[1];
                        } else {
                            x[u][0] = *reinterpret_cast<const P *>(xp + u * xstep);
                            v[u][0] = *reinterpret_cast<const P *>(vp + u * xstep);
                        }
#pragma unroll
                        for (int a = 0; a < KP; a += 2) {
                            const Vec<T> t = lds_vec<T>(wp + u * wstep + a * 8);
                            wd[u][a] = t.p[0];
                            wd[u][a + 1] = t.p[1];
                        }
                    }
                    xp += UU * xstep;
                    vp += UU * xstep;
                    wp += UU * wstep;
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
                    if (PROBE_OFF(32))  // bit 5: probe without role H's 2k accumulation FMAs
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
                for (; i + 2 <= nr; i += 2) rows_u(std::integral_constant<int, 2>());
                if (i < nr) rows_u(std::integral_constant<int, 1>());
#pragma unroll
                for (int c = 0; c < C2; ++c) chid += O::dsum(chi[c]);
            }
            __syncwarp(hmask);
            PROF_ADD(ht_loop, hb);
            if (lane == 0) {
                if (noprod == 1) {
                    // the column-owner warp that frees the stage last refills it
                    const unsigned old = atom_add_acqrel_a(freed_a + s * 4);
                    if ((old + 1u) % (unsigned)p.NW2 == 0u && nloc + stages < my_tiles) issue(nloc + stages);
                } else {
                    mbar_arrive_a(empty_a + s * 8);
                }
            }
            ++nloc;
            ready = done_next >= (unsigned)(pass_next + 1) * (unsigned)cpt;
#ifdef CNMF_TUNABLE
            if (!p.lookahead) ready = false;
#endif
            s = s_next;
            if (s == 0) ph ^= 1;
            pass = pass_next;
            if (SZ == 4 && ++since == p.flush_every) {
                flush();
                since = 0;
            }
        }
#ifdef CNMF_PROF
        if (blockIdx.x == 0 && lane == 0)
            printf("warp %2d role H: total %lld wait %lld rows %lld tiles %lld (ready on arrival %lld)\n", warp, clock64() - ht0,
                   ht_wait, ht_loop, hn, hn_ready);
#endif
        flush();
        }
        chid = warp_sum(chid);
        if (lane == 0) p.chi_part[blockIdx.x * p.NW2 + w2] = chid;
    }
}

}  // namespace cnmf
