// K1t -- the fused row-streaming pass for fp32, k <= 8, with every contraction of the iteration on the tensor cores
// (mma.sync.m16n8k8 TF32, three passes per product: hi*hi + hi*lo + lo*hi with x = hi + lo, hi = tf32(x), which keeps
// the products at fp32 accuracy -- the tolerance of this path is 1e-4 on iterates that K1 reproduces to 2e-6).
//
// Same pipeline and the same fragment chaining as mma_kernel.cuh (the fp64 kernel K1m): the accumulator layout of
// W H puts (row, column pair) on a lane, V*X and V*WH are formed there and go back in as A operands of (V*X) H^T,
// (V*WH) H^T with the column pair as the contracted index; role H is the transpose.  What differs:
//   * the tensor shape is 16 x 8 x 8: role W works on 16-row tiles against 8-column groups, role H on 16-column
//     groups against the two 8-row halves of a tile;
//   * H and W' are split into hi / lo once (H at kernel start, into two layouts; W' by the warp that finishes a
//     tile); V*X and V*WH are split where they are formed (one cvt + one subtraction per value);
//   * accumulators are fp32: the H-update partials are flushed into the fp64 partial block every `flush_every`
//     tiles, chi2 every tile, as in stream_kernel.cuh.
// Unlike fp64, where DMMA shares the FP64 pipe with DFMA, the TF32 tensor pipe is a unit of its own: 18 MMAs per
// 128 elements (2.25 cycles per element and SM sub-partition... 144 cycles per 128) replace 6k+7 FFMA per element
// that issue at 3 cycles per FFMA2 when all three operands are fresh (profiles/r1b_micro_regfile.txt).
#pragma once
#include "common.cuh"
#include "stream_kernel.cuh"

namespace cnmf {

struct Mma32Smem {
    int hrow_off, hp, ht_off, cp, stage_off, stage_bytes, pb, v_off, win_off, wd_off, wp_off, total, ng8, ng16;
};
// hp: floats per row of the row-major H copies; cp: columns covered (a multiple of 16); pb: bytes per row of X / V in a stage
__host__ __device__ inline Mma32Smem mma32_smem(int K, long long ld, int stages, int TW) {
    Mma32Smem s;
    auto up = [](long long v, long long a) { return (int)((v + a - 1) / a * a); };
    s.ng8 = (int)((ld + 7) / 8);
    s.ng16 = (int)((ld + 15) / 16);
    s.cp = s.ng16 * 16;
    s.hp = s.cp;
    while (s.hp % 32 != 8 && s.hp % 32 != 24) s.hp += 4;  // consecutive rows land 8 banks apart
    int pf = s.cp;
    while (pf % 32 != 8 && pf % 32 != 24) pf += 4;
    s.pb = pf * 4;
    s.hrow_off = up(40 * stages + 4, 128);                 // barriers and counters live below (as in mma_kernel.cuh)
    s.ht_off = up(s.hrow_off + 2LL * 8 * s.hp * 4, 128);   // H row-major, hi then lo
    s.stage_off = up(s.ht_off + 2LL * s.cp * 8 * 4, 128);  // H column-major with (k, k + 4) adjacent, hi then lo
    s.v_off = 16 * s.pb;
    s.win_off = 32 * s.pb;
    s.wd_off = s.win_off + up(16LL * K * 4, 128);
    s.wp_off = s.wd_off + 1024;  // W' hi [16][8], lo [16][8]
    s.stage_bytes = up(s.wp_off + (TW > 1 ? 2 : 0) * 1024, 128);
    s.total = s.stage_off + stages * s.stage_bytes;
    (void)K;
    return s;
}

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
// x = hi + lo.  The tensor core reads the top 19 bits of an fp32 register as TF32 and ignores the rest, so hi is x itself
// (used as trunc(x)) and lo = x - trunc(x), exact in fp32, goes in the same way: what is lost is 2^-20 of x.  Two
// instructions per value; cvt.rna.tf32.f32 compiles to four (compare with infinity, add, select, mask) plus the subtraction.
__device__ __forceinline__ void split_tf32(float x, uint32_t &hi, uint32_t &lo) {
    hi = __float_as_uint(x);
    lo = __float_as_uint(x - __uint_as_float(hi & 0xffffe000u));
}
// d += a * b with both operands split (three passes; the lo * lo term is below fp32 rounding)
__device__ __forceinline__ void mma3(float (&d)[4], const uint32_t (&ahi)[4], const uint32_t (&alo)[4], const uint32_t (&bhi)[2],
                                     const uint32_t (&blo)[2]) {
    mma_tf32(d, alo, bhi);
    mma_tf32(d, ahi, blo);
    mma_tf32(d, ahi, bhi);
}
__device__ __forceinline__ uint2 lds_u2(const unsigned char *p) { return *reinterpret_cast<const uint2 *>(p); }
__device__ __forceinline__ float2 lds_f2(const unsigned char *p) { return *reinterpret_cast<const float2 *>(p); }
__device__ __forceinline__ float lds_f1(const unsigned char *p) { return *reinterpret_cast<const float *>(p); }
__device__ __forceinline__ uint32_t lds_u1(const unsigned char *p) { return *reinterpret_cast<const uint32_t *>(p); }

// StreamParams fields used: as mma_kernel.cuh, plus flush_every (tiles between two flushes of the fp32 H partials).
template <int K, int GHMAX, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) mma32_kernel(const StreamParams<float> p) {
    static_assert(K >= 1 && K <= 8, "k <= 8");
    extern __shared__ __align__(128) unsigned char smem[];
    if (p.state->done) return;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, c = lane & 3;
    const int stages = p.stages, TW = p.crows;
    const long long ld = p.ld;
    const int rowb = (int)ld * 4;
    const Mma32Smem L = mma32_smem(K, ld, stages, TW);
    const int PB = L.pb, HPB = L.hp * 4, NG8 = L.ng8;
    uint64_t *full = reinterpret_cast<uint64_t *>(smem);
    uint64_t *wready = full + stages;
    uint64_t *empty = wready + stages;
    unsigned int *chunk_ctr = reinterpret_cast<unsigned int *>(empty + stages);
    volatile unsigned int *armed = chunk_ctr + 1;
    unsigned int *wcnt = chunk_ctr + 1 + stages;
    unsigned int *wpair = chunk_ctr + 1 + 2 * stages;
    unsigned char *Hrow = smem + L.hrow_off;           // [hi | lo][8][hp]
    unsigned char *Ht = smem + L.ht_off;               // [hi | lo][cp][8], component k at 2 (k & 3) + (k >> 2)
    const int hrow_lo = 8 * HPB, ht_lo = L.cp * 32;    // byte offsets of the lo copies
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
    // H -> shared, split, in both layouts; zero in the pad columns and in components K..7
    for (int i = tid; i < 8 * L.hp; i += blockDim.x) {
        const int a = i / L.hp, j = i - a * L.hp;
        const float h = (a < K && j < p.m) ? p.H[(long long)a * p.m + j] : 0.f;
        uint32_t hi, lo;
        split_tf32(h, hi, lo);
        reinterpret_cast<uint32_t *>(Hrow)[i] = hi;
        reinterpret_cast<uint32_t *>(Hrow + hrow_lo)[i] = lo;
        if (j < L.cp) {
            const int at = j * 8 + 2 * (a & 3) + (a >> 2);
            reinterpret_cast<uint32_t *>(Ht)[at] = hi;
            reinterpret_cast<uint32_t *>(Ht + ht_lo)[at] = lo;
        }
    }
    // the tail of every row slot (columns ld .. pitch) is never written by a copy: X = V = 0 there
    {
        const int padf = (PB - rowb) / 4;
        for (int i = tid; i < stages * 32 * padf; i += blockDim.x) {
            const int slot = i / padf, j = i - slot * padf;
            const int s = slot >> 5, r = slot & 31;
            reinterpret_cast<float *>(stage0 + s * sbytes + r * PB + rowb)[j] = 0.f;
        }
    }
    __syncthreads();

    const uint32_t full_a = smem_u32(full), wready_a = smem_u32(wready), empty_a = smem_u32(empty);
    const uint32_t wcnt_a = smem_u32(wcnt), wpair_a = smem_u32(wpair);
    const int R = 16;
    const int last_tile = p.n_tiles - 1;
    const int last_rows = (int)(p.n_rows - (long long)last_tile * R);
    const int my_tiles = p.n_tiles > (int)blockIdx.x ? (p.n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

    if (role == 0) {
        // ------------------------------------------------------------- producer (one warp, 33 copies per tile)
        int s = 0, ph = 0, pass = 0;
        for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
            mbar_wait_warp_a(empty_a + s * 8, ph ^ 1);
            const int rows = tile == last_tile ? last_rows : R;
            const uint32_t wb = (uint32_t)(rows * K * 4);
            const bool wbulk = (wb & 15u) == 0;
            const uint32_t st = smem_u32(stage0) + s * sbytes;
            if (lane == 0) mbar_arrive_expect_tx_a(full_a + s * 8, 2u * rows * rowb + (wbulk ? wb : 0u));
            __syncwarp();
            const long long r0 = (long long)tile * R;
            {
                const int r = lane & 15;
                if (r < rows) {
                    const float *src = (lane < 16 ? p.X : p.V) + (r0 + r) * ld;
                    bulk_g2s_a(st + (lane < 16 ? 0 : L.v_off) + r * PB, src, (uint32_t)rowb, full_a + s * 8);
                }
            }
            if (lane == 0 && wbulk) bulk_g2s_a(st + L.win_off, p.W + r0 * K, wb, full_a + s * 8);
            __syncwarp();
            if (lane == 0) armed[s] = (unsigned)pass + 1u;
            if (++s == stages) { s = 0; ph ^= 1; ++pass; }
        }
    } else if (role < 128) {
        // ------------------------------------------------- role W: row owners (a chunk = a column part of a 16-row tile)
        const int tw_sh = 31 - __clz(TW);
        const int GPC = (NG8 + TW - 1) / TW;
        const unsigned total_chunks = (unsigned)my_tiles << tw_sh;
        auto grab = [&]() {
            unsigned cc = 0;
            if (lane == 0) cc = atomicAdd(chunk_ctr, 1u);
            return __shfl_sync(0xffffffffu, cc, 0);
        };
        unsigned cq = grab();
        while (cq < total_chunks) {
            const unsigned cnext = grab();
            const int n = (int)(cq >> tw_sh);
            const int part = (int)(cq & (unsigned)(TW - 1));
            const int pass = (int)__umulhi((unsigned)n, p.stages_inv);
            const int s = n - pass * stages;
            const int tile = (int)blockIdx.x + n * (int)gridDim.x;
            while (!__all_sync(0xffffffffu, armed[s] > (unsigned)pass)) __nanosleep(p.update_w ? 32 : 400);
            mbar_wait_warp_a(full_a + s * 8, pass & 1);
            const int rows = tile == last_tile ? last_rows : R;
            unsigned char *st = stage0 + s * sbytes;
            const bool wbulk = ((rows * K * 4) & 15) == 0;
            const float *win = wbulk ? reinterpret_cast<const float *>(st + L.win_off) : p.W + (long long)tile * R * K;
            const bool v0 = g < rows, v1 = g + 8 < rows;  // this lane's two rows: g and g + 8
            float num[4] = {0.f, 0.f, 0.f, 0.f}, den[4] = {0.f, 0.f, 0.f, 0.f};
            if (p.update_w) {
                // A of W H: (row g, k c), (row g + 8, k c), (row g, k c + 4), (row g + 8, k c + 4)
                uint32_t whi[4], wlo[4];
                split_tf32((v0 && c < K) ? win[g * K + c] : 0.f, whi[0], wlo[0]);
                split_tf32((v1 && c < K) ? win[(g + 8) * K + c] : 0.f, whi[1], wlo[1]);
                split_tf32((v0 && c + 4 < K) ? win[g * K + (c + 4 < K ? c + 4 : 0)] : 0.f, whi[2], wlo[2]);
                split_tf32((v1 && c + 4 < K) ? win[(g + 8) * K + (c + 4 < K ? c + 4 : 0)] : 0.f, whi[3], wlo[3]);
                const int g0 = part * GPC, g1 = g0 + GPC < NG8 ? g0 + GPC : NG8;
                const unsigned char *xq = st + g * PB + g0 * 32 + c * 8;          // X[row g][8 grp + 2 c]; row g + 8 at + 8 PB
                const unsigned char *hB = Ht + (g0 * 8 + g) * 32 + c * 8;         // H[c][col], H[c + 4][col] of column 8 grp + g
                const unsigned char *hT = Hrow + g * HPB + g0 * 32 + c * 8;       // H[component g][8 grp + 2 c .. + 1]
#pragma unroll 2
                for (int grp = g0; grp < g1; ++grp) {
                    const float2 x0 = lds_f2(xq), x1 = lds_f2(xq + 8 * PB);
                    const float2 va = lds_f2(xq + L.v_off), vb = lds_f2(xq + L.v_off + 8 * PB);
                    const uint2 bh = lds_u2(hB), bl = lds_u2(hB + ht_lo);
                    const uint2 th = lds_u2(hT), tl = lds_u2(hT + hrow_lo);
                    const uint32_t bhi[2] = {bh.x, bh.y}, blo[2] = {bl.x, bl.y};
                    const uint32_t thi[2] = {th.x, th.y}, tlo[2] = {tl.x, tl.y};
                    float wh[4] = {0.f, 0.f, 0.f, 0.f};
                    mma3(wh, whi, wlo, bhi, blo);
                    // D: (row g, cols 2c, 2c + 1), (row g + 8, cols 2c, 2c + 1); as A the pair is the contracted index:
                    // a0 = (row g, 2c), a1 = (row g + 8, 2c), a2 = (row g, 2c + 1), a3 = (row g + 8, 2c + 1)
                    uint32_t xhi[4], xlo[4], yhi[4], ylo[4];
                    split_tf32(va.x * x0.x, xhi[0], xlo[0]);
                    split_tf32(vb.x * x1.x, xhi[1], xlo[1]);
                    split_tf32(va.y * x0.y, xhi[2], xlo[2]);
                    split_tf32(vb.y * x1.y, xhi[3], xlo[3]);
                    split_tf32(va.x * wh[0], yhi[0], ylo[0]);
                    split_tf32(vb.x * wh[2], yhi[1], ylo[1]);
                    split_tf32(va.y * wh[1], yhi[2], ylo[2]);
                    split_tf32(vb.y * wh[3], yhi[3], ylo[3]);
                    mma3(num, xhi, xlo, thi, tlo);
                    mma3(den, yhi, ylo, thi, tlo);
                    xq += 32;
                    hB += 8 * 32;
                    hT += 32;
                }
            }
            // column parts meet in shared memory in a fixed order (mma_kernel.cuh explains the two protocols)
            bool last = true;
            if (TW == 4) {
                const int pr = part >> 1;
                float *slot = reinterpret_cast<float *>(st + L.wp_off) + pr * 256 + lane * 8;
                if (!(part & 1)) {
                    if (p.update_w) {
                        *reinterpret_cast<float4 *>(slot) = make_float4(num[0], num[1], num[2], num[3]);
                        *reinterpret_cast<float4 *>(slot + 4) = make_float4(den[0], den[1], den[2], den[3]);
                    }
                    __syncwarp();
                    if (lane == 0) red_release_add_a(wpair_a + (pr * stages + s) * 4);
                    last = false;
                } else {
                    while (!__all_sync(0xffffffffu, lds_acquire_u32_a(wpair_a + (pr * stages + s) * 4) > (unsigned)pass))
                        __nanosleep(20);
                    if (p.update_w) {
                        const float4 n0 = *reinterpret_cast<const float4 *>(slot);
                        const float4 d0 = *reinterpret_cast<const float4 *>(slot + 4);
                        num[0] += n0.x; num[1] += n0.y; num[2] += n0.z; num[3] += n0.w;
                        den[0] += d0.x; den[1] += d0.y; den[2] += d0.z; den[3] += d0.w;
                        *reinterpret_cast<float4 *>(slot) = make_float4(num[0], num[1], num[2], num[3]);
                        *reinterpret_cast<float4 *>(slot + 4) = make_float4(den[0], den[1], den[2], den[3]);
                    }
                    __syncwarp();
                    unsigned old = 0;
                    if (lane == 0) old = atom_add_acqrel_a(wcnt_a + s * 4);
                    old = __shfl_sync(0xffffffffu, old, 0);
                    last = (old & 1u) != 0u;
                    if (last && p.update_w) {
                        const float *other = reinterpret_cast<const float *>(st + L.wp_off) + (pr ^ 1) * 256 + lane * 8;
                        const float4 n1 = *reinterpret_cast<const float4 *>(other);
                        const float4 d1 = *reinterpret_cast<const float4 *>(other + 4);
                        // pair 0 + pair 1 on whichever side this warp is
                        num[0] += n1.x; num[1] += n1.y; num[2] += n1.z; num[3] += n1.w;
                        den[0] += d1.x; den[1] += d1.y; den[2] += d1.z; den[3] += d1.w;
                    }
                }
            } else if (TW == 2) {
                float *slots = reinterpret_cast<float *>(st + L.wp_off) + lane * 8;
                if (p.update_w) {
                    *reinterpret_cast<float4 *>(slots + part * 256) = make_float4(num[0], num[1], num[2], num[3]);
                    *reinterpret_cast<float4 *>(slots + part * 256 + 4) = make_float4(den[0], den[1], den[2], den[3]);
                }
                __syncwarp();
                unsigned old = 0;
                if (lane == 0) old = atom_add_acqrel_a(wcnt_a + s * 4);
                old = __shfl_sync(0xffffffffu, old, 0);
                last = (old & 1u) != 0u;
                if (last && p.update_w) {
                    const float4 n0 = *reinterpret_cast<const float4 *>(slots), n1 = *reinterpret_cast<const float4 *>(slots + 256);
                    const float4 d0 = *reinterpret_cast<const float4 *>(slots + 4), d1 = *reinterpret_cast<const float4 *>(slots + 260);
                    num[0] = n0.x + n1.x; num[1] = n0.y + n1.y; num[2] = n0.z + n1.z; num[3] = n0.w + n1.w;
                    den[0] = d0.x + d1.x; den[1] = d0.y + d1.y; den[2] = d0.z + d1.z; den[3] = d0.w + d1.w;
                }
            }
            if (last) {
                // lane (g, c) holds num / den of W[row g][2c, 2c + 1] (entries 0, 1) and W[row g + 8][2c, 2c + 1] (2, 3)
                uint32_t *wdh = reinterpret_cast<uint32_t *>(st + L.wd_off), *wdl = wdh + 128;
                float wnew[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int r = g + 8 * (q >> 1), a = 2 * c + (q & 1);
                    float wn = 0.f;
                    if (r < rows && a < K) {
                        const float wc = win[r * K + a];
                        wn = p.update_w ? (wc * num[q]) / (den[q] > p.eps ? den[q] : p.eps) : wc;
                    }
                    uint32_t hi, lo;
                    split_tf32(wn, hi, lo);
                    wdh[r * 8 + a] = hi;
                    wdl[r * 8 + a] = lo;
                    wnew[q] = wn;
                }
                if (rows < R) {
                    // ragged last tile of the problem: the rows past its end hold stale bytes; role H needs X = V = 0
                    for (int r = rows; r < R; ++r)
                        for (int i = lane; i < PB / 4; i += 32) {
                            reinterpret_cast<float *>(st + r * PB)[i] = 0.f;
                            reinterpret_cast<float *>(st + L.v_off + r * PB)[i] = 0.f;
                        }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive_a(wready_a + s * 8);
                if (p.update_w) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int r = g + 8 * (q >> 1), a = 2 * c + (q & 1);
                        if (r < rows && a < K) p.W[((long long)tile * R + r) * K + a] = wnew[q];
                    }
                }
            }
            cq = cnext;
        }
    } else if (role - 128 < p.NW2) {
        // ---------------------------------------------- role H: column owners (whole 16-column groups per warp)
        const int w2 = role - 128;
        const int gs = p.hg0[w2], GH = p.hgn[w2];
        float nH[GHMAX][4], dH[GHMAX][4];
#pragma unroll
        for (int i = 0; i < GHMAX; ++i)
#pragma unroll
            for (int q = 0; q < 4; ++q) nH[i][q] = dH[i][q] = 0.f;
        double *blk = p.part + (long long)blockIdx.x * p.part_stride;
        bool first = true;
        auto flush = [&]() {
#pragma unroll
            for (int i = 0; i < GHMAX; ++i) {
                if (i < GH) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int col = (gs + i) * 16 + g + 8 * (q >> 1), a = 2 * c + (q & 1);
                        if (col < (int)ld && a < K) {
                            // every entry of the block belongs to one lane: the first flush of a launch overwrites what the last
                            // launch left, later ones add with a fire-and-forget reduction (a load-add-store chain per entry costs
                            // an L2 round trip each, 15 000 cycles per flush: 0.917 -> 0.816 ms per 10^6 rows at 8 x fewer flushes)
                            const long long in = (long long)a * ld + col, id = (long long)(K + a) * ld + col;
                            if (first) {
                                blk[in] = (double)nH[i][q];
                                blk[id] = (double)dH[i][q];
                            } else {
                                atomicAdd(blk + in, (double)nH[i][q]);
                                atomicAdd(blk + id, (double)dH[i][q]);
                            }
                        }
                        nH[i][q] = dH[i][q] = 0.f;
                    }
                }
            }
            first = false;
        };
        // byte offsets: X[row c (+ 4, + 8 h)][16 grp + g (+ 8)]
        const int xoff = c * PB + (gs * 16 + g) * 4;
        // W' fragments of a half h (rows 8 h ..): B of (W'H)^T = W'[8 h + rho(g)][c], [c + 4], rho(n) = n / 2 + 4 (n % 2);
        // B of the accumulations = W'[8 h + c][g], W'[8 h + c + 4][g]
        const int wBoff = L.wd_off + (((g >> 1) + 4 * (g & 1)) * 8 + c) * 4;
        const int wNoff = L.wd_off + (c * 8 + g) * 4;
        double chid = 0.0;
        int s = 0, ph = 0, since = 0;
        for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
            mbar_wait_warp_a(wready_a + s * 8, ph);
            const unsigned char *st = stage0 + s * sbytes;
            float chi = 0.f;
            // the (group, half) units of a tile as ONE straight-line block per group count (a block per unit leaves the
            // warp waiting on its own chain, as in mma_kernel.cuh)
            auto body = [&](auto g_tag) {
                constexpr int G = decltype(g_tag)::value;
                // W' fragments of both 8-row halves
                uint32_t wBh[2][2], wBl[2][2], wNh[2][2], wNl[2][2];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const unsigned char *wb = st + wBoff + h * 256, *wn = st + wNoff + h * 256;
                    wBh[h][0] = lds_u1(wb); wBh[h][1] = lds_u1(wb + 16); wBl[h][0] = lds_u1(wb + 512); wBl[h][1] = lds_u1(wb + 512 + 16);
                    wNh[h][0] = lds_u1(wn); wNh[h][1] = lds_u1(wn + 128); wNl[h][0] = lds_u1(wn + 512); wNl[h][1] = lds_u1(wn + 512 + 128);
                }
#pragma unroll
                for (int i = 0; i < G; ++i) {
                    // A of (W'H)^T: (col g, k c), (col g + 8, k c), (col g, k c + 4), (col g + 8, k c + 4): once per group
                    const unsigned char *ha = Ht + ((gs + i) * 16 + g) * 32 + c * 8;
                    const uint2 h0 = lds_u2(ha), h1 = lds_u2(ha + 8 * 32), l0 = lds_u2(ha + ht_lo), l1 = lds_u2(ha + ht_lo + 8 * 32);
                    const uint32_t ahi[4] = {h0.x, h1.x, h0.y, h1.y}, alo[4] = {l0.x, l1.x, l0.y, l1.y};
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const unsigned char *xq = st + xoff + h * 8 * PB + i * 64;
                        // (col g, row c), (col g, row c + 4), (col g + 8, row c), (col g + 8, row c + 4)
                        const float x0 = lds_f1(xq), x1 = lds_f1(xq + 4 * PB), x2 = lds_f1(xq + 32), x3 = lds_f1(xq + 4 * PB + 32);
                        const unsigned char *vq = xq + L.v_off;
                        const float v0 = lds_f1(vq), v1 = lds_f1(vq + 4 * PB), v2 = lds_f1(vq + 32), v3 = lds_f1(vq + 4 * PB + 32);
                        float wh[4] = {0.f, 0.f, 0.f, 0.f};
                        mma3(wh, ahi, alo, wBh[h], wBl[h]);
                        const float r0 = x0 - wh[0], r1 = x1 - wh[1], r2 = x2 - wh[2], r3 = x3 - wh[3];
                        chi = fmaf(v0 * r0, r0, chi);
                        chi = fmaf(v1 * r1, r1, chi);
                        chi = fmaf(v2 * r2, r2, chi);
                        chi = fmaf(v3 * r3, r3, chi);
                        // as A of the accumulations the row pair is the contracted index:
                        // a0 = (col g, row c), a1 = (col g + 8, row c), a2 = (col g, row c + 4), a3 = (col g + 8, row c + 4)
                        uint32_t xhi[4], xlo[4], yhi[4], ylo[4];
                        split_tf32(v0 * x0, xhi[0], xlo[0]);
                        split_tf32(v2 * x2, xhi[1], xlo[1]);
                        split_tf32(v1 * x1, xhi[2], xlo[2]);
                        split_tf32(v3 * x3, xhi[3], xlo[3]);
                        split_tf32(v0 * wh[0], yhi[0], ylo[0]);
                        split_tf32(v2 * wh[2], yhi[1], ylo[1]);
                        split_tf32(v1 * wh[1], yhi[2], ylo[2]);
                        split_tf32(v3 * wh[3], yhi[3], ylo[3]);
                        mma3(nH[i], xhi, xlo, wNh[h], wNl[h]);
                        mma3(dH[i], yhi, ylo, wNh[h], wNl[h]);
                    }
                }
            };
            if (GH == GHMAX) body(std::integral_constant<int, GHMAX>());
            else if (GHMAX > 1 && GH == GHMAX - 1) body(std::integral_constant<int, (GHMAX > 1 ? GHMAX - 1 : 1)>());
            else if (GHMAX > 2 && GH == GHMAX - 2) body(std::integral_constant<int, (GHMAX > 2 ? GHMAX - 2 : 1)>());
            else if (GHMAX > 3 && GH == GHMAX - 3) body(std::integral_constant<int, (GHMAX > 3 ? GHMAX - 3 : 1)>());
            chid += (double)chi;
            __syncwarp();
            if (lane == 0) mbar_arrive_a(empty_a + s * 8);
            if (++s == stages) { s = 0; ph ^= 1; }
            if (++since == p.flush_every) {
                flush();
                since = 0;
            }
        }
        flush();
        chid = warp_sum(chid);
        if (lane == 0) p.chi_part[blockIdx.x * p.NW2 + w2] = chid;
    }
}

}  // namespace cnmf
