// K1m -- the fused row-streaming pass for fp64, k <= 8, with every contraction of the iteration on the
// FP64 tensor path (DMMA, mma.sync.m8n8k4.f64).
//
// Same pipeline as stream_kernel.cuh (one persistent CTA per SM, a ring of row tiles filled by the TMA
// engine, row owners = role W, column owners = role H), different arithmetic.  DMMA runs on the same FP64
// pipe as DFMA (profiles/r1_micro_tensor_vs_fma.txt: 125.6 vs 114.8 MAC/ns/SM), so it buys no peak -- but
// stream_kernel does not sit at the peak: it holds the pipe at 54-57 % because three in-order warps per
// sub-partition issue ~110 instructions and 46 LSU cycles per 64 elements.  One
// DMMA does the work of eight warp-wide DFMAs and sums over the contracted index inside the instruction:
//
//   * role W: an 8-row tile against 8-column groups.  WH = W H (one DMMA for components 0..3; 4 and 5 are scalar
//     FMAs handed to it as C at k = 5, 6, a second DMMA at k = 7, 8), the lane that
//     receives (row, column pair) of WH multiplies by V and X held in the same layout, and V*X, V*WH
//     go straight back in as A operands of (V*X) H^T and (V*WH) H^T (4 DMMA): the sums over the columns
//     of a row -- the 8-lane shuffle reduction of stream_kernel -- happen inside the instruction and
//     lane (row, a) ends up holding num / den of W[row][a].  5 shared-memory loads per 64 elements
//     instead of 16.
//   * role H: transposed.  (W'H)^T = H^T W'^T puts the ROWS on the accumulator's paired index, so
//     V*X and V*W'H are A operands of (V*X)^T W' and (V*W'H)^T W' (4 DMMA) whose accumulators are the
//     H-update partials of 8 columns x 8 components: two registers pairs per lane and column group
//     instead of k x C2 per thread, and every lane of every warp is live (38 column groups over
//     8 warps) where stream_kernel's fifth column-owner warp runs 23 of 32 lanes.
//
// Cost: k = 6 pads to 8 in the four accumulations of each role: 10 DMMA + 22 DFMA-class instructions per 64
// elements = 204 pipe cycles per SM sub-partition against 172 for the scalar form.  The kernel pays where it holds
// the pipe well above the 54-57 % of stream_kernel (it reaches 70 %).
//
// Shared-memory layouts are chosen so that every fragment load is conflict-free: rows of a stage at a
// pitch of 8 or 24 (mod 32) words, H row-major at a pitch of 8 (mod 32) words, and the (free) mapping of
// fragment indices to rows / components permuted where a quarter-warp would otherwise hit one bank group
// twice (sigma below).  Rows are copied one bulk copy each (the pitch differs from the HBM pitch).
#pragma once
#include "common.cuh"
#include "stream_kernel.cuh"

namespace cnmf {

#ifdef CNMF_PROF
// diagnostics build: per-tile event times of CTA 0 (tools/prof_roles.py prints them with CNMF_TRACE=1):
// [tile][0] copies issued, [1] first row-owner warp saw them landed, [2] last row-owner part left its loop,
// [3] W' handed over, [4] first column-owner warp started the tile, [5] last one finished it (stage free)
#define CNMF_TRACE_TILES 1024
static __device__ unsigned long long g_mma_trace[CNMF_TRACE_TILES][6];
#define TRACE_SET(n, e) do { if (blockIdx.x == 0 && (n) < CNMF_TRACE_TILES && lane == 0) g_mma_trace[n][e] = (unsigned long long)clock64(); } while (0)
#define TRACE_MIN(n, e) do { if (blockIdx.x == 0 && (n) < CNMF_TRACE_TILES && lane == 0) atomicMin(&g_mma_trace[n][e], (unsigned long long)clock64()); } while (0)
#define TRACE_MAX(n, e) do { if (blockIdx.x == 0 && (n) < CNMF_TRACE_TILES && lane == 0) atomicMax(&g_mma_trace[n][e], (unsigned long long)clock64()); } while (0)
#else
#define TRACE_SET(n, e)
#define TRACE_MIN(n, e)
#define TRACE_MAX(n, e)
#endif

struct MmaSmem {
    int h_off, hp, stage_off, stage_bytes, pb, v_off, win_off, wd_off, wp_off, total, ng;
};
// hp: doubles per row of H in shared memory; pb: bytes per row of X / V in a stage
__host__ __device__ inline MmaSmem mma_smem(int K, long long ld, int stages, int TW) {
    MmaSmem s;
    auto up = [](long long v, long long a) { return (int)((v + a - 1) / a * a); };
    s.ng = (int)((ld + 7) / 8);
    s.hp = s.ng * 8;
    while (s.hp % 16 != 4) s.hp += 2;  // 2 * hp = 8 (mod 32) words
    s.pb = s.ng * 64;
    while ((s.pb / 4) % 32 != 8 && (s.pb / 4) % 32 != 24) s.pb += 16;
    s.h_off = up(40 * stages + 4, 128);  // barriers (3 per stage) and counters (chunk queue, 4 per stage) live below
    s.stage_off = up(s.h_off + 8LL * s.hp * 8, 128);
    s.v_off = 8 * s.pb;
    s.win_off = 16 * s.pb;
    s.wd_off = s.win_off + up(8LL * K * 8, 128);
    s.wp_off = s.wd_off + 512;
    s.stage_bytes = up(s.wp_off + (TW > 1 ? 2 : 0) * 1024, 128);  // slots of 32 lanes x 4 sums: one per part (TW = 2) or per pair of parts
    s.total = s.stage_off + stages * s.stage_bytes;
    return s;
}

// volatile: the DMMAs of a block are issued in the order written (the orders below keep dependent ones apart;
// left to itself the compiler finishes one column group before it starts the next)
__device__ __forceinline__ void dmma(double (&d)[2], double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
        : "+d"(d[0]), "+d"(d[1])
        : "d"(a), "d"(b));
}
__device__ __forceinline__ double lds_f64(const unsigned char *p) { return *reinterpret_cast<const double *>(p); }

// sigma: fragment index 0..7 -> row of the tile (role W) / component of H (role W's H^T fragment).  The
// two indices of a quarter-warp (2q, 2q + 1) must land two pitches apart.
__device__ __forceinline__ int mma_sigma(int i) { return (i & 4) | ((i & 1) << 1) | ((i >> 1) & 1); }

// StreamParams fields used: X V W H part chi_part state n_rows ld part_stride m n_tiles NW2 stages
// stages_inv update_w eps role[] gate_ns prefetch; crows carries TW (column parts of a tile in role W, 1 / 2 / 4),
// lookahead the suspend-time hint (ns) of the mbarrier waits.
template <int K, int GHMAX, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) mma_kernel(const StreamParams<double> p) {
    static_assert(K >= 1 && K <= 8, "k <= 8");
    constexpr bool K2 = K > 6;  // second k-step of the W H products on the tensor path (components 4..7)
    // k = 5, 6: a second DMMA would be 3/4 resp. 1/2 padding (16 pipe cycles); components 4 and 5 are two
    // resp. four scalar FMAs per lane instead (4 / 8 cycles), formed FIRST and handed to the DMMA as its C
    constexpr int KSC = (K == 5 || K == 6) ? K - 4 : 0;
    extern __shared__ __align__(128) unsigned char smem[];
    if (p.state->done) return;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m8 = lane >> 2, c4 = lane & 3;
    const int stages = p.stages, TW = p.crows;
    const long long ld = p.ld;
    const int rowb = (int)ld * 8;
    const MmaSmem L = mma_smem(K, ld, stages, TW);
    const int PB = L.pb, HPB = L.hp * 8, NG = L.ng;
    uint64_t *full = reinterpret_cast<uint64_t *>(smem);
    uint64_t *wready = full + stages;
    uint64_t *empty = wready + stages;
    unsigned int *chunk_ctr = reinterpret_cast<unsigned int *>(empty + stages);
    volatile unsigned int *armed = chunk_ctr + 1;  // [stages]: how often a stage was armed (row owners claim chunks ahead of the copies)
    unsigned int *wcnt = chunk_ctr + 1 + stages;   // [stages]: parts (TW = 2) / pairs of parts (TW = 4) that have left their sums
    unsigned int *wpair = chunk_ctr + 1 + 2 * stages;  // [2][stages]: even parts of a pair that have left their sums (TW = 4)
    unsigned char *Hs = smem + L.h_off;
    unsigned char *const stage0 = smem + L.stage_off;
    const int sbytes = L.stage_bytes;
    const int role = p.role[warp];

    if (tid == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&wready[s], 1);
            mbar_init(&empty[s], p.NW2);
            armed[s] = 0u;
            wcnt[s] = 0u;
            wpair[s] = 0u;
            wpair[stages + s] = 0u;
        }
        *chunk_ctr = 0u;
        mbar_fence_init();
        if (blockIdx.x == 0) p.state->k1_count += 1;
    }
    // H -> shared, row-major at pitch hp, zero in the pad columns and in rows K..7
    for (int i = tid; i < 8 * L.hp; i += blockDim.x) {
        const int a = i / L.hp, j = i - a * L.hp;
        reinterpret_cast<double *>(Hs)[i] = (a < K && j < p.m) ? p.H[(long long)a * p.m + j] : 0.0;
    }
    // the tail of every row slot (columns ld .. pitch) is never written by a copy: X = V = 0 there
    {
        const int padd = (PB - rowb) / 8;
        for (int i = tid; i < stages * 16 * padd; i += blockDim.x) {
            const int slot = i / padd, j = i - slot * padd;
            const int s = slot >> 4, r = slot & 15;
            reinterpret_cast<double *>(stage0 + s * sbytes + r * PB + rowb)[j] = 0.0;
        }
    }
    __syncthreads();

    const uint32_t full_a = smem_u32(full), wready_a = smem_u32(wready), empty_a = smem_u32(empty);
    const uint32_t wcnt_a = smem_u32(wcnt), wpair_a = smem_u32(wpair);
    const int R = 8;
    const int last_tile = p.n_tiles - 1;
    const int last_rows = (int)(p.n_rows - (long long)last_tile * R);
    const int my_tiles = p.n_tiles > (int)blockIdx.x ? (p.n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

    if (role == 0) {
        // ------------------------------------------------------------- producer (one warp, 17 copies per tile)
        int s = 0, ph = 0, pass = 0;
        PROF_DECL(pw = 0, pn = 0);
        PROF_T(pt0);
        for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
            if (p.prefetch > 0 && lane < 2) {
                // pull a later tile of this CTA towards L2 (its rows are contiguous in HBM): the stage it goes to is
                // still in use, but the copy that follows will then take an L2 round trip instead of a DRAM one
                const long long tpf = (long long)tile + (long long)p.prefetch * gridDim.x;
                if (tpf < p.n_tiles) {
                    const int prow = tpf == last_tile ? last_rows : R;
                    bulk_prefetch_l2((lane ? p.V : p.X) + tpf * R * ld, (uint32_t)(prow * rowb));
                }
            }
            PROF_T(pa);
            mbar_wait_warp_hint_a(empty_a + s * 8, ph ^ 1, p.lookahead);
            PROF_ADD(pw, pa);
#ifdef CNMF_PROF
            ++pn;
#endif
            const int rows = tile == last_tile ? last_rows : R;
            const uint32_t wb = (uint32_t)(rows * K * 8);
            const bool wbulk = (wb & 15u) == 0;
            const uint32_t st = smem_u32(stage0) + s * sbytes;
            if (lane == 0) mbar_arrive_expect_tx_a(full_a + s * 8, 2u * rows * rowb + (wbulk ? wb : 0u));
            __syncwarp();
            const long long r0 = (long long)tile * R;
            if (lane < 16) {
                const int r = lane & 7;
                if (r < rows) {
                    const double *src = (lane < 8 ? p.X : p.V) + (r0 + r) * ld;
                    bulk_g2s_a(st + (lane < 8 ? 0 : L.v_off) + r * PB, src, (uint32_t)rowb, full_a + s * 8);
                }
            } else if (lane == 16 && wbulk) {
                bulk_g2s_a(st + L.win_off, p.W + r0 * K, wb, full_a + s * 8);
            }
            __syncwarp();
#ifdef CNMF_PROF
            if (lane == 0) reinterpret_cast<volatile long long *>(smem + L.total)[s] = clock64();
            TRACE_SET(pass * stages + s, 0);
#endif
            if (lane == 0) armed[s] = (unsigned)pass + 1u;
            if (++s == stages) { s = 0; ph ^= 1; ++pass; }
        }
#ifdef CNMF_PROF
        if (blockIdx.x == 0 && lane == 0) printf("warp %2d producer: total %lld wait-empty %lld tiles %lld\n", warp, clock64() - pt0, pw, pn);
#endif
    } else if (role < 128) {
        // ------------------------------------------------- role W: row owners (a chunk = a column part of a tile)
        const int tw_sh = 31 - __clz(TW);
        const int GPC = (NG + TW - 1) / TW;
        const unsigned total_chunks = (unsigned)my_tiles << tw_sh;
        const int row = mma_sigma(m8);     // the row of the tile this lane's A / D fragments belong to
        const int acomp = mma_sigma(m8);   // the component of H this lane feeds as B operand of num / den
        auto grab = [&]() {
            unsigned c = 0;
            if (lane == 0) c = atomicAdd(chunk_ctr, 1u);
            return __shfl_sync(0xffffffffu, c, 0);
        };
        PROF_DECL(wt_gate = 0, wt_full = 0, wt_loop = 0, wt_tail = 0, wn = 0, wt_lat = 0, wn_lat = 0);
        PROF_T(wt0);
        unsigned c = grab();
        while (c < total_chunks) {
            const unsigned cnext = grab();
            PROF_T(wa);
            const int n = (int)(c >> tw_sh);
            const int part = (int)(c & (unsigned)(TW - 1));
            const int pass = (int)__umulhi((unsigned)n, p.stages_inv);
            const int s = n - pass * stages;
            const int tile = (int)blockIdx.x + n * (int)gridDim.x;
            while (!__all_sync(0xffffffffu, armed[s] > (unsigned)pass)) __nanosleep(p.update_w ? p.gate_ns : 400);
            PROF_ADD(wt_gate, wa);
            PROF_T(wb_);
            mbar_wait_warp_hint_a(full_a + s * 8, pass & 1, p.lookahead);
            PROF_ADD(wt_full, wb_);
            TRACE_MIN(n, 1);
#ifdef CNMF_PROF
            if (clock64() - wb_ > 64) {  // the copy had not landed when this warp came for it: issue -> landed, an upper bound
                wt_lat += clock64() - reinterpret_cast<volatile long long *>(smem + L.total)[s];
                ++wn_lat;
            }
#endif
            PROF_T(wc_);
#ifdef CNMF_PROF
            ++wn;
#endif
            const int rows = tile == last_tile ? last_rows : R;
            unsigned char *st = stage0 + s * sbytes;
            const bool wbulk = ((rows * K * 8) & 15) == 0;
            const double *win = wbulk ? reinterpret_cast<const double *>(st + L.win_off) : p.W + (long long)tile * R * K;
            const bool rvalid = row < rows;
            double num[2] = {0.0, 0.0}, den[2] = {0.0, 0.0};
            if (p.update_w) {
                const int g0 = part * GPC, g1 = g0 + GPC < NG ? g0 + GPC : NG;
                const double aW0 = (rvalid && c4 < K) ? win[row * K + c4] : 0.0;
                const double aW1 = (K2 && rvalid && c4 + 4 < K) ? win[row * K + (c4 + 4 < K ? c4 + 4 : 0)] : 0.0;
                const double w4 = (KSC >= 1 && rvalid) ? win[row * K + (K > 4 ? 4 : 0)] : 0.0;
                const double w5 = (KSC >= 2 && rvalid) ? win[row * K + (K > 5 ? 5 : 0)] : 0.0;
                const unsigned char *h45 = Hs + 4 * HPB + g0 * 64 + c4 * 16;  // H[4][columns of this lane's pair]; H[5] one pitch on
                auto wh_init = [&](double (&wh)[2], int o) {
                    if (KSC >= 1) {
                        const Vec<double> h4 = lds_vec<double>(h45 + o);
                        wh[0] = w4 * h4.p[0];
                        wh[1] = w4 * h4.p[1];
                        if (KSC >= 2) {
                            const Vec<double> h5 = lds_vec<double>(h45 + o + HPB);
                            wh[0] = fma(w5, h5.p[0], wh[0]);
                            wh[1] = fma(w5, h5.p[1], wh[1]);
                        }
                    } else {
                        wh[0] = wh[1] = 0.0;
                    }
                };
                const unsigned char *xq = st + row * PB + g0 * 64 + c4 * 16;
                const unsigned char *vq = xq + L.v_off;
                const unsigned char *hT = Hs + acomp * HPB + g0 * 64 + c4 * 16;
                const unsigned char *hB = Hs + c4 * HPB + (g0 * 8 + m8) * 8;
                double numb[2] = {0.0, 0.0}, denb[2] = {0.0, 0.0};
                auto one = [&](int o) {
                    const Vec<double> x = lds_vec<double>(xq + o);
                    const Vec<double> v = lds_vec<double>(vq + o);
                    const Vec<double> t = lds_vec<double>(hT + o);
                    double wh[2];
                    wh_init(wh, o);
                    dmma(wh, aW0, lds_f64(hB + o));
                    if (K2) dmma(wh, aW1, lds_f64(hB + o + 4 * HPB));
                    dmma(num, v.p[0] * x.p[0], t.p[0]);
                    dmma(den, v.p[0] * wh[0], t.p[0]);
                    dmma(num, v.p[1] * x.p[1], t.p[1]);
                    dmma(den, v.p[1] * wh[1], t.p[1]);
                };
                // Two column groups per pass, skewed by one pass: the eight accumulations of the previous pair are
                // issued behind the W H products of the current one (they fill the latency of its result and of
                // the V * WH products); pvx / pvw / pt carry V*X, V*WH and the H^T fragments of the pending pair.
                double pvx[4], pvw[4], pt[4];
                auto front = [&](double (&wa)[2], double (&wb)[2], Vec<double> &xa, Vec<double> &xb, Vec<double> &va,
                                 Vec<double> &vb, Vec<double> &ta, Vec<double> &tb) {
                    xa = lds_vec<double>(xq);
                    xb = lds_vec<double>(xq + 64);
                    va = lds_vec<double>(vq);
                    vb = lds_vec<double>(vq + 64);
                    ta = lds_vec<double>(hT);
                    tb = lds_vec<double>(hT + 64);
                    wh_init(wa, 0);
                    wh_init(wb, 64);
                    dmma(wa, aW0, lds_f64(hB));
                    dmma(wb, aW0, lds_f64(hB + 64));
                    if (K2) {
                        dmma(wa, aW1, lds_f64(hB + 4 * HPB));
                        dmma(wb, aW1, lds_f64(hB + 64 + 4 * HPB));
                    }
                    xq += 128;
                    vq += 128;
                    hT += 128;
                    hB += 128;
                    h45 += 128;
                };
                auto back = [&](const double (&wa)[2], const double (&wb)[2], const Vec<double> &xa, const Vec<double> &xb,
                                const Vec<double> &va, const Vec<double> &vb, const Vec<double> &ta, const Vec<double> &tb) {
                    pvx[0] = va.p[0] * xa.p[0]; pvx[1] = va.p[1] * xa.p[1];
                    pvx[2] = vb.p[0] * xb.p[0]; pvx[3] = vb.p[1] * xb.p[1];
                    pvw[0] = va.p[0] * wa[0]; pvw[1] = va.p[1] * wa[1];
                    pvw[2] = vb.p[0] * wb[0]; pvw[3] = vb.p[1] * wb[1];
                    pt[0] = ta.p[0]; pt[1] = ta.p[1]; pt[2] = tb.p[0]; pt[3] = tb.p[1];
                };
                auto flush_pair = [&]() {
                    dmma(num, pvx[0], pt[0]);
                    dmma(numb, pvx[2], pt[2]);
                    dmma(den, pvw[0], pt[0]);
                    dmma(denb, pvw[2], pt[2]);
                    dmma(num, pvx[1], pt[1]);
                    dmma(numb, pvx[3], pt[3]);
                    dmma(den, pvw[1], pt[1]);
                    dmma(denb, pvw[3], pt[3]);
                };
                int g = g0;
                if (g + 2 <= g1) {
                    {
                        double wa[2], wb[2];
                        Vec<double> xa, xb, va, vb, ta, tb;
                        front(wa, wb, xa, xb, va, vb, ta, tb);
                        back(wa, wb, xa, xb, va, vb, ta, tb);
                    }
                    g += 2;
#pragma unroll 1
                    for (; g + 2 <= g1; g += 2) {
                        double wa[2], wb[2];
                        Vec<double> xa, xb, va, vb, ta, tb;
                        front(wa, wb, xa, xb, va, vb, ta, tb);
                        flush_pair();
                        back(wa, wb, xa, xb, va, vb, ta, tb);
                    }
                    flush_pair();
                }
                if (g < g1) one(0);
                num[0] += numb[0];
                num[1] += numb[1];
                den[0] += denb[0];
                den[1] += denb[1];
            }
            PROF_ADD(wt_loop, wc_);
            TRACE_MAX(n, 2);
            PROF_T(wd_);
            // W' of the rows: with one part per tile straight from the registers; otherwise the column parts
            // of the tile meet in shared memory and the part that arrives last sums them in part order
            bool last = true;
            if (TW == 4) {
                // Four column parts, two slots: the even part of a pair leaves its sums in the pair's slot; the odd part
                // waits for that (chunks are claimed in part order, so it seldom has to), adds its own -- two operands,
                // so the sum does not depend on who was first -- and leaves the pair's sum there; of the two odd parts
                // the one that arrives second adds pair 0 + pair 1 and finishes the rows.  Every sum has a fixed order.
                const int pr = part >> 1;
                double *slot = reinterpret_cast<double *>(st + L.wp_off) + pr * 128 + lane * 4;
                if (!(part & 1)) {
                    if (p.update_w) {
                        *reinterpret_cast<double2 *>(slot) = make_double2(num[0], num[1]);
                        *reinterpret_cast<double2 *>(slot + 2) = make_double2(den[0], den[1]);
                    }
                    __syncwarp();
                    if (lane == 0) red_release_add_a(wpair_a + (pr * stages + s) * 4);
                    last = false;
                } else {
                    while (!__all_sync(0xffffffffu, lds_acquire_u32_a(wpair_a + (pr * stages + s) * 4) > (unsigned)pass))
                        __nanosleep(20);
                    if (p.update_w) {
                        const double2 n0 = *reinterpret_cast<const double2 *>(slot);
                        const double2 d0 = *reinterpret_cast<const double2 *>(slot + 2);
                        num[0] += n0.x; num[1] += n0.y; den[0] += d0.x; den[1] += d0.y;
                        *reinterpret_cast<double2 *>(slot) = make_double2(num[0], num[1]);
                        *reinterpret_cast<double2 *>(slot + 2) = make_double2(den[0], den[1]);
                    }
                    __syncwarp();
                    unsigned old = 0;
                    if (lane == 0) old = atom_add_acqrel_a(wcnt_a + s * 4);
                    old = __shfl_sync(0xffffffffu, old, 0);
                    last = (old & 1u) != 0u;
                    if (last && p.update_w) {
                        const double *other = reinterpret_cast<const double *>(st + L.wp_off) + (pr ^ 1) * 128 + lane * 4;
                        const double2 n1 = *reinterpret_cast<const double2 *>(other);
                        const double2 d1 = *reinterpret_cast<const double2 *>(other + 2);
                        num[0] += n1.x; num[1] += n1.y; den[0] += d1.x; den[1] += d1.y;
                    }
                }
            } else if (TW > 1) {
                double *slots = reinterpret_cast<double *>(st + L.wp_off) + lane * 4;
                if (p.update_w) {
                    *reinterpret_cast<double2 *>(slots + part * 128) = make_double2(num[0], num[1]);
                    *reinterpret_cast<double2 *>(slots + part * 128 + 2) = make_double2(den[0], den[1]);
                }
                __syncwarp();
                unsigned old = 0;
                if (lane == 0) old = atom_add_acqrel_a(wcnt_a + s * 4);
                old = __shfl_sync(0xffffffffu, old, 0);
                last = (old & (unsigned)(TW - 1)) == (unsigned)(TW - 1);
                if (last && p.update_w) {
                    const double2 n0 = *reinterpret_cast<const double2 *>(slots);
                    const double2 d0 = *reinterpret_cast<const double2 *>(slots + 2);
                    num[0] = n0.x; num[1] = n0.y; den[0] = d0.x; den[1] = d0.y;
                    for (int q = 1; q < TW; ++q) {
                        const double2 nq = *reinterpret_cast<const double2 *>(slots + q * 128);
                        const double2 dq = *reinterpret_cast<const double2 *>(slots + q * 128 + 2);
                        num[0] += nq.x; num[1] += nq.y; den[0] += dq.x; den[1] += dq.y;
                    }
                }
            }
            if (last) {
                // lane (m8, c4) holds num / den of W[row][sigma(2 c4 + j)], j = 0, 1
                double *wd = reinterpret_cast<double *>(st + L.wd_off) + row * 8;
                const long long grow = (long long)tile * R + row;
                double wnew[2];
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const int a = mma_sigma(2 * c4 + j);
                    double wn = 0.0;
                    if (rvalid && a < K) {
                        const double wc = win[row * K + a];
                        if (p.update_w) {
                            const double dj = den[j] > p.eps ? den[j] : p.eps;
                            wn = ratio_div(wc * num[j], dj);
                        } else {
                            wn = wc;
                        }
                    }
                    wd[a] = wn;
                    wnew[j] = wn;
                }
                if (rows < R) {
                    // ragged last tile of the problem: the rows past its end hold stale bytes; role H needs X = V = 0
                    for (int r = rows; r < R; ++r)
                        for (int i = lane; i < PB / 8; i += 32) {
                            reinterpret_cast<double *>(st + r * PB)[i] = 0.0;
                            reinterpret_cast<double *>(st + L.v_off + r * PB)[i] = 0.0;
                        }
                }
                __syncwarp();
                TRACE_SET(n, 3);
                if (lane == 0) mbar_arrive_a(wready_a + s * 8);  // release: the column owners may read W' (and the zero-filled rows)
                // W' to HBM only now: the store has nothing to do with the hand-over and must not sit in front of it
                if (p.update_w && rvalid) {
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        const int a = mma_sigma(2 * c4 + j);
                        if (a < K) p.W[grow * K + a] = wnew[j];
                    }
                }
            }
            PROF_ADD(wt_tail, wd_);
            c = cnext;
        }
#ifdef CNMF_PROF
        if (blockIdx.x == 0 && lane == 0)
            printf("warp %2d role W: total %lld gate %lld full %lld loop %lld tail %lld chunks %lld; copy issue -> seen landed %lld cycles avg over %lld waits\n",
                   warp, clock64() - wt0, wt_gate, wt_full, wt_loop, wt_tail, wn, wn_lat ? wt_lat / wn_lat : 0, wn_lat);
#endif
    } else if (role - 128 < p.NW2) {
        // ---------------------------------------------- role H: column owners (whole 8-column groups per warp)
        const int w2 = role - 128;
        const int gs = p.hg0[w2], GH = p.hgn[w2];  // dealt by the host (api.cu): more groups where fewer warps share the sub-partition
        double nH[GHMAX][2], dH[GHMAX][2];
        // (the fragments of H a group needs come from shared memory each time: held in registers they cost 40 and the
        // kernel spilled, profiles/r2b_mma_role_timers_first.txt)
#pragma unroll
        for (int i = 0; i < GHMAX; ++i) nH[i][0] = nH[i][1] = dH[i][0] = dH[i][1] = 0.0;
        const unsigned char *hAp = Hs + c4 * HPB + (gs * 8 + m8) * 8;
        // byte offsets inside a stage: x of rows c4 (t = 0) and c4 + 4 (t = 1), column gs * 8 + m8
        const int xoff = c4 * PB + (gs * 8 + m8) * 8;
        // W' fragments: B of (W'H)^T is W'[row rho(m8)][c4 + 4 s], rho(n) = n / 2 + 4 (n % 2); B of the
        // accumulations is W'[row c4 + 4 t][component m8]
        const int wBoff = L.wd_off + ((m8 >> 1) + 4 * (m8 & 1)) * 64 + c4 * 8;
        const int wNoff = L.wd_off + c4 * 64 + m8 * 8;
        double chi = 0.0;
        int s = 0, ph = 0, pass = 0;
        PROF_DECL(ht_wait = 0, ht_loop = 0, hn = 0, hn_ready = 0);
        PROF_T(ht0);
        for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
            PROF_T(ha);
#ifdef CNMF_PROF
            ++hn;
#endif
            mbar_wait_warp_hint_a(wready_a + s * 8, ph, p.lookahead);
            PROF_ADD(ht_wait, ha);
            TRACE_MIN(pass * stages + s, 4);
            PROF_T(hb);
            const int s_next = (s + 1 == stages) ? 0 : s + 1;
            const int pass_next = (s + 1 == stages) ? pass + 1 : pass;
            const unsigned char *st = stage0 + s * sbytes;
            const double wB0 = lds_f64(st + wBoff), wB1 = K2 ? lds_f64(st + wBoff + 32) : 0.0;
            const double wN0 = lds_f64(st + wNoff), wN1 = lds_f64(st + wNoff + 256);
            // W'[row c4 + 4 j][4], [5] for the scalar components
            const double w40 = KSC >= 1 ? lds_f64(st + L.wd_off + c4 * 64 + 32) : 0.0;
            const double w41 = KSC >= 1 ? lds_f64(st + L.wd_off + (c4 + 4) * 64 + 32) : 0.0;
            const double w50 = KSC >= 2 ? lds_f64(st + L.wd_off + c4 * 64 + 40) : 0.0;
            const double w51 = KSC >= 2 ? lds_f64(st + L.wd_off + (c4 + 4) * 64 + 40) : 0.0;
            const unsigned char *hSp = Hs + 4 * HPB + (gs * 8 + m8) * 8;
            auto wh_init = [&](int i, double (&wh)[2]) {
                if (KSC >= 1) {
                    const double h4 = lds_f64(hSp + i * 64);
                    wh[0] = w40 * h4;
                    wh[1] = w41 * h4;
                    if (KSC >= 2) {
                        const double h5 = lds_f64(hSp + i * 64 + HPB);
                        wh[0] = fma(w50, h5, wh[0]);
                        wh[1] = fma(w51, h5, wh[1]);
                    }
                } else {
                    wh[0] = wh[1] = 0.0;
                }
            };
            const unsigned char *xp = st + xoff;
            // one column group: (W'H)^T, the element-wise part on the lane that owns (column, row pair), the
            // four accumulations.  Two groups run as one straight-line block so that the chains of one fill
            // the latencies of the other (a column-owner warp is the critical path of a tile).
            auto loads = [&](int i, double (&xv)[4], double (&h)[2]) {
                xv[0] = lds_f64(xp + i * 64);
                xv[1] = lds_f64(xp + i * 64 + 4 * PB);
                xv[2] = lds_f64(xp + i * 64 + L.v_off);
                xv[3] = lds_f64(xp + i * 64 + L.v_off + 4 * PB);
                h[0] = lds_f64(hAp + i * 64);
                h[1] = !K2 ? 0.0 : lds_f64(hAp + i * 64 + 4 * HPB);
            };
            // element-wise part on the lane that owns (column, row pair): V*X, V*W'H, chi2
            auto mid = [&](const double (&wh)[2], const double (&xv)[4], double (&o)[4]) {
                o[0] = xv[2] * xv[0];
                o[1] = xv[3] * xv[1];
                const double r0 = xv[0] - wh[0], r1 = xv[1] - wh[1];
                o[2] = xv[2] * wh[0];
                o[3] = xv[3] * wh[1];
                chi = fma(xv[2] * r0, r0, chi);
                chi = fma(xv[3] * r1, r1, chi);
            };
            auto acc = [&](int i, const double (&o)[4]) {
                dmma(nH[i], o[0], wN0);
                dmma(dH[i], o[2], wN0);
                dmma(nH[i], o[1], wN1);
                dmma(dH[i], o[3], wN1);
            };
            // G groups as ONE straight-line block, skewed by one group: the four accumulations of group i - 1 are
            // issued behind the W'H product of group i and fill the latency of its result and of the element-wise
            // chain that follows (a block per group, or a pair the compiler finishes one after the other, leaves
            // the warp waiting on its own chain -- and a column-owner warp is the critical path of a tile)
            auto body = [&](auto g_tag) {
                constexpr int G = decltype(g_tag)::value;
                double op[4];
                {
                    double w[2], x[4], h[2];
                    loads(0, x, h);
                    wh_init(0, w);
                    dmma(w, h[0], wB0);
                    if (K2) dmma(w, h[1], wB1);
                    mid(w, x, op);
                }
#pragma unroll
                for (int i = 1; i < G; ++i) {
                    double w[2], x[4], h[2], oc[4];
                    loads(i, x, h);
                    wh_init(i, w);
                    dmma(w, h[0], wB0);
                    if (K2) dmma(w, h[1], wB1);
                    acc(i - 1, op);
                    mid(w, x, oc);
#pragma unroll
                    for (int q = 0; q < 4; ++q) op[q] = oc[q];
                }
                acc(G - 1, op);
            };
            if (GH == GHMAX) {
                body(std::integral_constant<int, GHMAX>());
            } else if (GHMAX > 1 && GH == GHMAX - 1) {
                body(std::integral_constant<int, (GHMAX > 1 ? GHMAX - 1 : 1)>());
            } else if (GHMAX > 2 && GH == GHMAX - 2) {
                body(std::integral_constant<int, (GHMAX > 2 ? GHMAX - 2 : 1)>());
            } else {
                // any other count (narrow problems): a block per group
#pragma unroll
                for (int i = 0; i < GHMAX; ++i) {
                    if (i < GH) {
                        double w[2], x[4], h[2], o[4];
                        loads(i, x, h);
                        wh_init(i, w);
                        dmma(w, h[0], wB0);
                        if (K2) dmma(w, h[1], wB1);
                        mid(w, x, o);
                        acc(i, o);
                    }
                }
            }
            __syncwarp();
            PROF_ADD(ht_loop, hb);
            TRACE_MAX(pass * stages + s, 5);
            if (lane == 0) mbar_arrive_a(empty_a + s * 8);
            s = s_next;
            if (s == 0) ph ^= 1;
            pass = pass_next;
        }
#ifdef CNMF_PROF
        if (blockIdx.x == 0 && lane == 0)
            printf("warp %2d role H: total %lld wait %lld groups %lld tiles %lld (ready on arrival %lld)\n", warp, clock64() - ht0, ht_wait,
                   ht_loop, hn, hn_ready);
#endif
        // lane (m8, c4) of group i holds the partials of column (gs + i) * 8 + m8, components 2 c4 + j
        double *blk = p.part + (long long)blockIdx.x * p.part_stride;
#pragma unroll
        for (int i = 0; i < GHMAX; ++i) {
            const int col = (gs + i) * 8 + m8;
            if (i < GH && col < (int)ld) {
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const int a = 2 * c4 + j;
                    if (a < K) {
                        blk[(long long)a * ld + col] = nH[i][j];
                        blk[(long long)(K + a) * ld + col] = dH[i][j];
                    }
                }
            }
        }
        chi = warp_sum(chi);
        if (lane == 0) p.chi_part[blockIdx.x * p.NW2 + w2] = chi;
    }
}

}  // namespace cnmf
