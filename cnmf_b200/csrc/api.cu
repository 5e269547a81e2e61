// C ABI of libcnmf_b200.so (include/cnmf_b200.h): plans, the iteration loop
// (CUDA-graph captured), chi2, the host-buffer entry point.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <string>
#include <mutex>
#include <vector>

#include "../../include/cnmf_b200.h"
#include "comm.h"
#include "common.cuh"
#include "launch.h"
#include "stream_kernel.cuh"
#include "mma_kernel.cuh"
#include "mma32_kernel.cuh"

namespace cnmf {

// ---------------------------------------------------------------- errors
static thread_local char g_err[512] = "";
void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
static int cuda_fail(cudaError_t e, const char *what) {
    set_error("%s: CUDA error %d (%s)", what, (int)e, cudaGetErrorString(e));
    return 1000 + (int)e;
}
#define CU(call)                                        \
    do {                                                \
        cudaError_t e__ = (call);                       \
        if (e__ != cudaSuccess) return cuda_fail(e__, #call); \
    } while (0)

static std::atomic<long long> g_launches{0};
static thread_local bool g_capturing = false;  // launches recorded into a graph are counted per replay
void count_launch() {
    if (!g_capturing) g_launches.fetch_add(1, std::memory_order_relaxed);
}

// ---------------------------------------------------------------- K1 dispatch
#define DECL_K(TN, K) int launch_stream_##TN##_k##K(const StreamConfig &, const StreamArgs &, cudaStream_t);
#define DECL_ALL(TN) \
    DECL_K(TN, 1) DECL_K(TN, 2) DECL_K(TN, 3) DECL_K(TN, 4) DECL_K(TN, 5) DECL_K(TN, 6) DECL_K(TN, 7) DECL_K(TN, 8) \
    DECL_K(TN, 9) DECL_K(TN, 10) DECL_K(TN, 11) DECL_K(TN, 12) DECL_K(TN, 13) DECL_K(TN, 14) DECL_K(TN, 15) DECL_K(TN, 16)
DECL_ALL(f32)
DECL_ALL(f64)
typedef int (*stream_fn)(const StreamConfig &, const StreamArgs &, cudaStream_t);
#define ROW(TN)                                                                                                  \
    {launch_stream_##TN##_k1,  launch_stream_##TN##_k2,  launch_stream_##TN##_k3,  launch_stream_##TN##_k4,  \
     launch_stream_##TN##_k5,  launch_stream_##TN##_k6,  launch_stream_##TN##_k7,  launch_stream_##TN##_k8,  \
     launch_stream_##TN##_k9,  launch_stream_##TN##_k10, launch_stream_##TN##_k11, launch_stream_##TN##_k12, \
     launch_stream_##TN##_k13, launch_stream_##TN##_k14, launch_stream_##TN##_k15, launch_stream_##TN##_k16}
static const stream_fn g_stream[2][CNMF_MAX_K] = {ROW(f32), ROW(f64)};
// fp64 tensor-path kernel (mma_kernel.cuh), k <= 8
#define DECL_M(K) int launch_mma_k##K(const StreamConfig &, const StreamArgs &, cudaStream_t);
DECL_M(1) DECL_M(2) DECL_M(3) DECL_M(4) DECL_M(5) DECL_M(6) DECL_M(7) DECL_M(8)
static const stream_fn g_mma[8] = {launch_mma_k1, launch_mma_k2, launch_mma_k3, launch_mma_k4,
                                   launch_mma_k5, launch_mma_k6, launch_mma_k7, launch_mma_k8};

#define DECL_T(K) int launch_mma32_k##K(const StreamConfig &, const StreamArgs &, cudaStream_t);
DECL_T(1) DECL_T(2) DECL_T(3) DECL_T(4) DECL_T(5) DECL_T(6) DECL_T(7) DECL_T(8)
static const stream_fn g_mma32[8] = {launch_mma32_k1, launch_mma32_k2, launch_mma32_k3, launch_mma32_k4,
                                     launch_mma32_k5, launch_mma32_k6, launch_mma32_k7, launch_mma32_k8};

int launch_stream(const StreamConfig &c, const StreamArgs &a, cudaStream_t st) {
    if (c.mma == 5) return g_mma32[c.k - 1](c, a, st);
    if (c.mma) return g_mma[c.k - 1](c, a, st);
    return g_stream[c.dtype][c.k - 1](c, a, st);
}
bool stream_variant_exists(int dtype, int k, int G, int C2) {
    return (dtype == 0 || dtype == 1) && k >= 1 && k <= CNMF_MAX_K && (G == 8 || G == 16) && (C2 == 1 || C2 == 2);
}

// ---------------------------------------------------------------- config
static int env_tune(const char *key, int dflt) {
    // CNMF_TUNE="G=8,C2=1,RP1=2,R=16,NW1=6,S2=2,ST=5,MAXW=12,DEAL=1,FL=64,GRAPH=1": sweep knobs, see DESIGN.md.
    // Read by the tuning / diagnostics builds only (cnmf_b200/build.py --tune, --prof): the product library
    // ignores the variable, so a stray setting cannot select an untested variant or a timing probe.
#ifndef CNMF_TUNABLE
    (void)key;
    return dflt;
#else
    const char *s = getenv("CNMF_TUNE");
    if (!s) return dflt;
    std::string str(s), k = std::string(key) + "=";
    size_t pos = 0;
    while ((pos = str.find(k, pos)) != std::string::npos) {
        if (pos == 0 || str[pos - 1] == ',') return atoi(str.c_str() + pos + k.size());
        pos += k.size();
    }
    return dflt;
#endif
}

// Deal the warp roles so that the FMA-pipe work of the four SM sub-partitions (a warp
// lives on sub-partition warp % 4) is as even as the warp counts allow: the column
// owners and the row owners issue different numbers of FMAs per tile row, and the
// sub-partition with the most work sets the pace of the whole CTA.
static void deal_roles(int nw1, int nw2, double wcost, const double *hcost, unsigned char *role, int h_high,
                       bool producer = true) {
    const int total = (producer ? 1 : 0) + nw1 + nw2;
    double load[4] = {0, 0, 0, 0};
    int used[4] = {0, 0, 0, 0}, cap[4];
    for (int q = 0; q < 4; ++q) cap[q] = (total - q + 3) / 4;
    // items: column owners (their own costs), row owners, the producer; heaviest first, each onto
    // the least-loaded sub-partition that still has a free warp slot
    struct Item { double cost; int kind, idx; } it[32];
    int n = 0;
    for (int i = 0; i < nw2; ++i) it[n++] = {hcost[i], 2, i};
    for (int i = 0; i < nw1; ++i) it[n++] = {wcost, 1, i};
    if (producer) it[n++] = {0.0, 0, 0};
    for (int i = 0; i < n; ++i)
        for (int j = i + 1; j < n; ++j)
            if (it[j].cost > it[i].cost) { Item t = it[i]; it[i] = it[j]; it[j] = t; }
    int kind[32], idx[32];
    for (int i = 0; i < 32; ++i) kind[i] = -1;
    for (int i = 0; i < n; ++i) {
        int best = -1;
        for (int q = 0; q < 4; ++q)
            if (used[q] < cap[q] && (best < 0 || load[q] < load[best])) best = q;
        load[best] += it[i].cost;
        const int slot = best + 4 * used[best]++;
        kind[slot] = it[i].kind;
        idx[slot] = it[i].idx;
    }
    // Within a sub-partition the order of the warps is free.  h_high = 1 gives the column owners
    // the highest warp indices of their sub-partition, -1 the lowest (no effect measured).
    if (h_high != 0) {
        for (int q = 0; q < 4; ++q)
            for (int w = q; w < total; w += 4)
                for (int x = w + 4; x < total; x += 4)
                    if ((h_high > 0) ? kind[x] < kind[w] : kind[x] > kind[w]) {
                        int t = kind[w]; kind[w] = kind[x]; kind[x] = t;
                        t = idx[w]; idx[w] = idx[x]; idx[x] = t;
                    }
    }
    memset(role, 0xff, 32);
    int iw = 0;
    for (int w = 0; w < total; ++w) {
        if (kind[w] == 0) role[w] = 0;
        else if (kind[w] == 1) role[w] = (unsigned char)(1 + iw++);
        else if (kind[w] == 2) role[w] = (unsigned char)(128 + idx[w]);
    }
}

// The fp64 tensor-path kernel (mma_kernel.cuh): 8-row tiles, column-owner warps that own whole 8-column
// groups, row-owner warps that share a tile by column parts.  Chosen where it measured faster than the
// scalar stream kernel on a B200 (profiles/r2b_mma_shapes.txt, r2d_mma_policy.txt): k = 7, 8 from 12 column groups up
// (2.7-3 x at 301 columns; a tie at 81), k = 6 from 18 groups (>= 137 columns; 8-10 % at the bench size) and k = 5 from 36
// (2 % at 301 columns) while five stages fit (<= 320 columns); k <= 4 pads too much of the 8-wide tensor shape and stays scalar.  CNMF_MMA=0 forces the scalar kernel,
// CNMF_MMA=3 / 1 force the 16-warp / 12-warp form of the tensor-path kernel wherever it fits (tests run
// every setting); CNMF_MMA_TW = column parts of a tile in role W (1, 2, 4).
static bool choose_config_mma(const cnmf_desc_t &d, int num_sms, int max_smem, StreamConfig *out) {
    if (d.dtype != CNMF_F64 || d.k > 8) return false;
    const int ng = (int)((d.ld + 7) / 8);
    const char *e = getenv("CNMF_MMA");
    int mode = e ? atoi(e) : -1;
    const bool by_policy = mode < 0;
    if (by_policy) mode = ((d.k >= 7 && ng >= 12) || (d.k == 6 && ng >= 18) || (d.k == 5 && ng >= 36)) ? 3 : 0;  // the measured policy
    if (mode == 0) return false;
#ifdef CNMF_TUNABLE
    if (mode != 1 && mode != 4) mode = 3;
#else
    if (mode != 1) mode = 3;
#endif
    StreamConfig c;
    memset(&c, 0, sizeof(c));
    c.dtype = d.dtype;
    c.k = d.k;
    c.mma = mode;
    c.G = 32;
    c.C2 = 1;
    c.RP1 = 1;
    c.R = 8;
    c.S2 = 1;
    c.NT = 1;
    // accumulators of a column-owner warp: up to 6 column groups in the 16-warp form (5 on average at 320 columns), 10 in the 12-warp form
    const int nw2max = mode >= 3 ? 8 : 4, gpw = mode >= 3 ? kMmaGroupsPerWarp / 2 : kMmaGroupsPerWarp;
    const int gcap = mode >= 3 ? kMmaGroupsPerWarp / 2 + 1 : kMmaGroupsPerWarp;
    c.NW2 = ng < nw2max ? ng : nw2max;
    (void)gpw;
    if (ng > c.NW2 * gcap) return false;  // wider than the column owners' accumulators (384 columns in the 16-warp form)
    const int maxw = mode == 4 ? 17 : (mode == 3 ? 16 : 12);
    c.NW1 = maxw - 1 - c.NW2;
    const char *tw = getenv("CNMF_MMA_TW");
    // column parts of a tile in role W: four from 32 column groups up (301 columns: -1.4 % at k = 6, -4.5 % at k = 7, 8
    // against two), two down to 12 groups (150 / 184 / 200 columns: four parts measured 3-5 % slower), one below
    c.crows = tw ? atoi(tw) : (ng >= 32 ? 4 : (ng >= 12 ? 2 : 1));
    if (c.crows != 1 && c.crows != 2 && c.crows != 4) c.crows = 2;
    while (c.crows > 1 && ng < 2 * c.crows) c.crows >>= 1;
    c.NW1 = env_tune("NW1", c.NW1) < c.NW1 && env_tune("NW1", c.NW1) >= 1 ? env_tune("NW1", c.NW1) : c.NW1;
    int stages = env_tune("ST", 8);
    if (stages > 8) stages = 8;
    for (; stages >= 2; --stages)
        if (mma_smem(d.k, d.ld, stages, c.crows).total <= max_smem) break;
    if (stages < 3) return false;
    // k = 5, 6 pay only with the full ring: at 383 columns (three 49 KB stages) the scalar kernel is 13 % faster
    if (by_policy && d.k <= 6 && stages < 5) return false;
    c.stages = stages;
    c.smem_bytes = mma_smem(d.k, d.ld, stages, c.crows).total;
    c.TPS = c.NW2 * 32;
    c.threads = 32 * (1 + c.NW1 + c.NW2);
    // column owners on warps 0 .. NW2-1 (spread over the SM sub-partitions), then the row owners, the producer last
    memset(c.role, 0xff, 32);
    for (int i = 0; i < c.NW2; ++i) c.role[i] = (unsigned char)(128 + i);
    for (int i = 0; i < c.NW1; ++i) c.role[c.NW2 + i] = (unsigned char)(1 + i);
    c.role[c.NW2 + c.NW1] = 0;
    // Column groups per column-owner warp.  A tile's stage is free when the SLOWEST column-owner warp is done with it,
    // and a warp's pace is set by how many warps share its SM sub-partition (warp % 4): deal the groups in proportion
    // to 1 / (compute warps on the sub-partition), largest remainders first, so that the warps finish a tile together.
    {
        int per_sp[4] = {0, 0, 0, 0};
        for (int w = 0; w < c.NW2 + c.NW1; ++w) per_sp[w & 3]++;
        double speed[8], tot = 0;
        for (int i = 0; i < c.NW2; ++i) tot += (speed[i] = 1.0 / per_sp[i & 3]);
        int n[8], left = ng;
        double frac[8];
        const int even = env_tune("HGEVEN", 0);
        for (int i = 0; i < c.NW2; ++i) {
            const double want = even ? (double)ng / c.NW2 : ng * speed[i] / tot;
            n[i] = (int)want;
            if (n[i] > gcap) n[i] = gcap;
            frac[i] = want - n[i];
            left -= n[i];
        }
        while (left > 0) {
            int best = -1;
            for (int i = 0; i < c.NW2; ++i)
                if (n[i] < gcap && (best < 0 || frac[i] > frac[best])) best = i;
            if (best < 0) return false;
            n[best]++;
            frac[best] -= 1.0;
            --left;
        }
        const int hg = env_tune("HG", 0);  // tuning builds: the counts as decimal digits, warp 0 first (HG=44564456)
        if (hg > 0) {
            int dgt[8], m = hg, sum = 0;
            for (int i = c.NW2 - 1; i >= 0; --i) { dgt[i] = m % 10; m /= 10; sum += dgt[i]; }
            if (sum == ng) for (int i = 0; i < c.NW2; ++i) n[i] = dgt[i];
        }
        int g0 = 0;
        for (int i = 0; i < c.NW2; ++i) {
            c.hg0[i] = (unsigned char)g0;
            c.hgn[i] = (unsigned char)n[i];
            g0 += n[i];
        }
    }
    const long long n_tiles = (d.n_rows + c.R - 1) / c.R;
    int sms = num_sms;
    if (d.sm_limit > 0 && d.sm_limit < sms) sms = d.sm_limit;
    c.grid = (int)(n_tiles < sms ? n_tiles : sms);
    if (c.grid < 1) c.grid = 1;
    c.flush_every = 1 << 30;
    c.chi_slots = c.NW2;
    c.gate_ns = env_tune("GATENS", 32);
    c.prefetch = env_tune("PF", 0);
    c.lookahead = env_tune("HINT", 0);  // suspend-time hint of the mbarrier waits, ns
    *out = c;
    return true;
}

// The fp32 tensor-path kernel (mma32_kernel.cuh, TF32 x 3): 16-row tiles, column-owner warps that own whole 16-column
// groups.  Parity green, measured 6 % (k = 6, 301 columns) to 14 % (k = 8, 184 columns) SLOWER than the FFMA2 stream
// kernel (profiles/r2c_mma32_tf32.txt), so it exists in the tuning builds only: CNMF_MMA32=1 selects it there.
static bool choose_config_mma32(const cnmf_desc_t &d, int num_sms, int max_smem, StreamConfig *out) {
#ifndef CNMF_TUNABLE
    (void)d; (void)num_sms; (void)max_smem; (void)out;
    return false;
#else
    if (d.dtype != CNMF_F32 || d.k > 8) return false;
    const char *e = getenv("CNMF_MMA32");
    int mode = e ? atoi(e) : 0;
    if (mode <= 0) return false;
    StreamConfig c;
    memset(&c, 0, sizeof(c));
    c.dtype = d.dtype;
    c.k = d.k;
    c.mma = 5;
    c.G = 32;
    c.C2 = 1;
    c.RP1 = 1;
    c.R = 16;
    c.S2 = 1;
    c.NT = 1;
    const int ng8 = (int)((d.ld + 7) / 8), ng16 = (int)((d.ld + 15) / 16);
    c.NW2 = ng16 < 8 ? ng16 : 8;
    if (ng16 > c.NW2 * kMma32GroupsPerWarp) return false;
    c.NW1 = 15 - c.NW2;
    if (c.NW1 > 8) c.NW1 = 8;
    const char *tw = getenv("CNMF_MMA_TW");
    c.crows = tw ? atoi(tw) : (ng8 >= 32 ? 4 : (ng8 >= 12 ? 2 : 1));
    if (c.crows != 1 && c.crows != 2 && c.crows != 4) c.crows = 2;
    while (c.crows > 1 && ng8 < 2 * c.crows) c.crows >>= 1;
    {
        const int nw1 = env_tune("NW1", c.NW1);
        if (nw1 >= 1 && nw1 < c.NW1) c.NW1 = nw1;
    }
    int stages = env_tune("ST", 8);
    if (stages > 8) stages = 8;
    for (; stages >= 2; --stages)
        if (mma32_smem(d.k, d.ld, stages, c.crows).total <= max_smem) break;
    if (stages < 3) return false;
    c.stages = stages;
    c.smem_bytes = mma32_smem(d.k, d.ld, stages, c.crows).total;
    c.TPS = c.NW2 * 32;
    c.threads = 32 * (1 + c.NW1 + c.NW2);
    memset(c.role, 0xff, 32);
    for (int i = 0; i < c.NW2; ++i) c.role[i] = (unsigned char)(128 + i);
    for (int i = 0; i < c.NW1; ++i) c.role[c.NW2 + i] = (unsigned char)(1 + i);
    c.role[c.NW2 + c.NW1] = 0;
    {
        // 16-column groups per column-owner warp: as even as they go
        const int base = ng16 / c.NW2, rem = ng16 - base * c.NW2;
        int g0 = 0;
        for (int i = 0; i < c.NW2; ++i) {
            const int n = base + (i >= c.NW2 - rem ? 1 : 0);
            c.hg0[i] = (unsigned char)g0;
            c.hgn[i] = (unsigned char)n;
            g0 += n;
        }
    }
    const long long n_tiles = (d.n_rows + c.R - 1) / c.R;
    int sms = num_sms;
    if (d.sm_limit > 0 && d.sm_limit < sms) sms = d.sm_limit;
    c.grid = (int)(n_tiles < sms ? n_tiles : sms);
    if (c.grid < 1) c.grid = 1;
    c.flush_every = env_tune("FL", 32);  // tiles between two flushes of the fp32 H partials into the fp64 block (512 rows)
    if (c.flush_every < 1) c.flush_every = 1;
    c.chi_slots = c.NW2;
    *out = c;
    return true;
#endif
}

static int choose_config(const cnmf_desc_t &d, int num_sms, int max_smem, StreamConfig *out) {
    if (choose_config_mma(d, num_sms, max_smem, out)) return 0;
    if (choose_config_mma32(d, num_sms, max_smem, out)) return 0;
    // Defaults come from the sweeps recorded in profiles/ (round 1): few, fat threads --
    // 12 warps per CTA (<= 384 threads, 168 registers), role H with two packs per thread,
    // role W with 8 lanes per row and two rows per lane.  Every knob can be overridden
    // through CNMF_TUNE for sweeps.
    StreamConfig c;
    memset(&c, 0, sizeof(c));
    c.dtype = d.dtype;
    c.k = d.k;
    const bool f32 = d.dtype == CNMF_F32;
    const int sz = f32 ? 4 : 8;
    const int pw = f32 ? 2 : 1;
    const long long ld = d.ld;
    const int packs = (int)(ld / pw);
    const int ldv = (int)(ld * sz / 16);
    c.G = env_tune("G", 8);
    if (c.G != 8 && c.G != 16) c.G = 8;
    // fp32: two rows per row-owner lane halve the H-vector loads (k = 8, 184 bands: 1.61 -> 0.98 ms,
    // profiles/r1_sweeps_v4.txt); fp64 cannot buffer the 8-row chunks that needs (47 rows fit)
    c.RP1 = env_tune("RP1", (f32 && d.k <= 8) ? 2 : 1) == 2 ? 2 : 1;
    // register-resident role W (stream_kernel NI > 0): a row-owner warp takes whole rows, H of its
    // packs in registers; exists where ceil(packs / 32) has a compiled instance (launch.h)
    {
        const int ni = (packs + 31) / 32;
        const int nvk = d.k <= 1 ? 1 : (d.k <= 2 ? 2 : (d.k <= 4 ? 4 : (d.k <= 8 ? 8 : 16)));
        if (env_tune("WREG", 0) && stream_wreg_ok(f32, d.k, ni)) {
            c.wreg = ni;
            c.G = 32;
            c.RP1 = 1;
            int crw = env_tune("CR", 2);
            if (crw > 32 / nvk) crw = 32 / nvk;
            if (crw < 1) crw = 1;
            while (crw & (crw - 1)) --crw;  // power of two, so that it divides the tile rows
            c.crows = crw;
        }
    }
    int c2 = env_tune("C2", d.k <= 6 ? 2 : 1);
    if (c2 != 1 && c2 != 2) c2 = 1;
    // CTA-size cap: the fat variants exist under the 384-thread cap, C2 = 2 also under 512
    int maxw = env_tune("MAXW", 12);
    if (maxw > stream_max_threads(d.k) / 32) maxw = stream_max_threads(d.k) / 32;
    if ((c.RP1 == 2 || c.wreg) && maxw > 12) maxw = 12;
    if (c2 == 2 && maxw > 16) maxw = 16;
    if (maxw < 3) maxw = 3;
    c.noprod = env_tune("NOPROD", 0);  // 1: last column owner to free a stage refills it; 2: an idle lane is the producer
    if (c.noprod < 0 || c.noprod > 2) c.noprod = 0;
    const int pw_ = c.noprod ? 0 : 1;  // warps spent on a producer
    const int cr = c.wreg ? c.crows : (32 / c.G) * c.RP1;  // rows per role-W chunk
    // rows per tile: the largest power-of-two multiple of the chunk (up to four chunks) that still leaves a
    // ring of four stages.  (Until round 1b: the largest whose X+V stage stays under 40 KB -- the same
    // choice except for stages of 40..52 KB, where the bigger tile measured 2-10 % faster: 184 bands fp32,
    // k = 4 / 6 / 7 / 8: +1.5 / 2.7 / 5.7 / 8.5 %, profiles/r1b_config3_sweep.txt.)
    int R = env_tune("R", 0);
    if (R <= 0) {
        R = c.wreg ? (cr > 32 ? cr : 32) : 4 * cr;
        auto ring = [&](int rows) {
            int st = 16;
            for (; st >= 2; --st)
                if (stream_smem(d.k, ld, sz, rows, st).total <= max_smem) break;
            return st;
        };
        while (R > cr && ring(R) < 4) R >>= 1;
    }
    if (R % cr) R = (R / cr + 1) * cr;
    {
        // chunks per tile: a power of two (the kernel maps a chunk index to its tile with a shift)
        int q = R / cr;
        while (q & (q - 1)) --q;
        R = q * cr;
    }
    c.R = R;
    int stages = env_tune("ST", 0);
    if (stages <= 0) stages = 16;
    for (; stages >= 2; --stages) {
        StreamSmem L = stream_smem(d.k, ld, sz, R, stages);
        if (L.total <= max_smem) break;
    }
    if (stages < 2) {
        set_error("shared memory too small for m=%d k=%d", d.m, d.k);
        return CNMF_ERR_UNSUPPORTED;
    }
    c.stages = stages;
    c.smem_bytes = stream_smem(d.k, ld, sz, R, stages).total;
    // role H: S2 row streams of TPS threads each
    if ((packs + c2 - 1) / c2 > (maxw - 1 - pw_) * 32 && c2 == 1) c2 = 2;
    c.C2 = c2;
    c.TPS = (packs + c2 - 1) / c2;
    if (c.TPS > (maxw - 1 - pw_) * 32) {
        set_error("m=%d is too wide for k=%d in this build (role-H threads %d > %d)", d.m, d.k, c.TPS, (maxw - 1 - pw_) * 32);
        return CNMF_ERR_UNSUPPORTED;
    }
    // FMA-pipe instructions per tile row: all column owners together / all row owners together
    const double h_total = (3.0 * d.k + 5) * ((c.TPS + 31) / 32) * c2;  // per stream-row: warps x packs
    const double w_total = c.wreg ? (3.0 * d.k + 2) * c.wreg + 2.5 * d.k + 6  // + the 32-lane reduction and the W' arithmetic
                                  : (3.0 * d.k + 2) * 2.0 * ((ldv + c.G - 1) / c.G) * c.G / 32.0;
    int s2 = env_tune("S2", 0);
    if (s2 <= 0) {
        // streams: as many as keep the column owners' share of the warps near their share of the work
        const double want = (maxw - pw_) * h_total / (h_total + w_total);
        s2 = 1;
        for (int s = 1; s <= R; ++s)
            if (R % s == 0 && (s * c.TPS + 31) / 32 <= want + 0.5 && (s * c.TPS + 31) / 32 <= maxw - 1 - pw_) s2 = s;
    }
    if (R % s2) {
        set_error("S2=%d does not divide the tile rows %d", s2, R);
        return CNMF_ERR_UNSUPPORTED;
    }
    c.S2 = s2;
    c.NW2 = (s2 * c.TPS + 31) / 32;
    // a last warp that is at most three quarters full is run as NT = 2 row-interleaved copies
    // (its cost lands on two sub-partitions instead of one)
    // -- but only when the row owners can spare the warp: measured at M = 301, 12 warps: k = 4 gains
    // 8 %, k = 6 loses 3 % (profiles/r1_sweeps_v4.txt, NT sweep)
    const int last_lanes = s2 * c.TPS - 32 * (c.NW2 - 1);
    const int nw1_after = maxw - 1 - pw_ - c.NW2;
    const bool spare = nw1_after >= 1 && w_total / nw1_after <= 0.8 * (3.0 * d.k + 5) * c2 / s2;
    c.NT = env_tune("NT", (last_lanes <= 24 && R % (2 * s2) == 0 && spare) ? 2 : 1);
    if (c.NT < 1 || R % (c.NT * s2) != 0 || c.NW2 + c.NT - 1 > maxw - 1 - pw_) c.NT = 1;
    c.NW2 += c.NT - 1;
    if (c.noprod == 2 && last_lanes > 31) c.noprod = 1;  // no idle lane to carry the producer loop
    int nw1 = env_tune("NW1", 0);
    if (nw1 <= 0) nw1 = maxw - pw_ - c.NW2;
    if (nw1 > maxw - pw_ - c.NW2) nw1 = maxw - pw_ - c.NW2;
    if (nw1 < 1) {
        set_error("no warps left for the W role (k=%d, R=%d, G=%d, role-H warps %d of %d)", d.k, R, c.G, c.NW2, maxw);
        return CNMF_ERR_UNSUPPORTED;
    }
    c.NW1 = nw1;
    c.threads = 32 * (pw_ + nw1 + c.NW2);
    if ((c.threads > 512 && c.C2 == 2) || (c.threads > 384 && c.RP1 == 2) || c.threads > stream_max_threads(d.k)) {
        set_error("no compiled stream-kernel instance for %d threads with C2=%d RP1=%d", c.threads, c.C2, c.RP1);
        return CNMF_ERR_UNSUPPORTED;
    }
    const int deal = env_tune("DEAL", 1);  // 0 plain order, 1 balanced, 2 balanced + column owners high, 3 ... low
    if (deal) {
        double hc[32];
        const double per_warp = (3.0 * d.k + 5) * c2 / s2;  // a full-row-set column-owner warp, per tile row
        for (int i = 0; i < c.NW2; ++i) hc[i] = i >= c.NW2 - c.NT ? per_warp / c.NT : per_warp;
        deal_roles(nw1, c.NW2, w_total / nw1, hc, c.role, deal == 2 ? 1 : (deal == 3 ? -1 : 0), !c.noprod);
    }
    else {
        memset(c.role, 0xff, 32);  // plain order: producer, row owners, column owners
        if (pw_) c.role[0] = 0;
        for (int i = 0; i < nw1; ++i) c.role[pw_ + i] = (unsigned char)(1 + i);
        for (int i = 0; i < c.NW2; ++i) c.role[pw_ + nw1 + i] = (unsigned char)(128 + i);
    }
    const long long n_tiles = (d.n_rows + R - 1) / R;
    int sms = num_sms;
    if (d.sm_limit > 0 && d.sm_limit < sms) sms = d.sm_limit;
    c.grid = (int)(n_tiles < sms ? n_tiles : sms);
    if (c.grid < 1) c.grid = 1;
    int fl = env_tune("FL", 0);
    if (fl <= 0) fl = (512 * s2 + R - 1) / R;
    c.flush_every = fl < 1 ? 1 : fl;
#ifdef CNMF_PROF
    c.debug = env_tune("DBG", 0);  // probe bits of the diagnostics build; the product kernels never read them
#endif
    c.prefetch = env_tune("PF", 0);
    c.gate_ns = env_tune("GATENS", 32);
    c.lookahead = env_tune("LOOKAHEAD", 1);
    c.chi_slots = c.NW2;
    *out = c;
    return 0;
}

// ---------------------------------------------------------------- plan
struct GraphKey {
    const void *X, *V, *W, *H, *hist;
    int n, update;
    double tol;
    cudaStream_t st;
    bool operator==(const GraphKey &o) const {
        return X == o.X && V == o.V && W == o.W && H == o.H && hist == o.hist && n == o.n && update == o.update &&
               tol == o.tol && st == o.st;
    }
};

}  // namespace cnmf

struct cnmf_plan {
    cnmf_desc_t d;
    cnmf::StreamConfig cfg;
    int device = 0;
    int num_sms = 0;
    int batched = 0;
    int batched_threads = 0, batched_smem = 0;
    int res_cs = 0, res_rows = 0, res_smem = 0;  // resident cluster kernel (resident.cu): CTAs per problem, rows per CTA
    long long n_tiles = 0;
    long long part_stride = 0;
    int n_entries = 0;  // 2 * k * ld
    int nblk = 0, n_chi = 0;
    double *part = nullptr, *chi_part = nullptr, *red = nullptr;
    cnmf::State *state = nullptr;
    void *scratch = nullptr;         // the one device allocation part / chi_part / red / state point into
    cnmf::State *h_state = nullptr;  // pinned (from the pool below)
    cudaStream_t own_stream = nullptr;
    cudaEvent_t ev_in = nullptr, ev_out = nullptr;
    bool use_graph = true;
    bool have_graph = false;
    cnmf::GraphKey gkey;
    cudaGraphExec_t gexec = nullptr;
    long long scratch_bytes = 0;
    std::vector<cudaEvent_t> tev;  // start/stop pairs of the timed fused-kernel launches
    int tev_used = 0;
    bool timing = false;
};

namespace cnmf {

static double default_eps(int dtype) { return dtype == CNMF_F32 ? 1e-12 : 1e-16; }

// Pinned State blocks are pooled: cudaMallocHost costs about a millisecond, more than a small solve.
static std::mutex g_pin_mu;
static std::vector<State *> g_pin_free;
static State *pinned_state_get() {
    {
        std::lock_guard<std::mutex> lk(g_pin_mu);
        if (!g_pin_free.empty()) {
            State *s = g_pin_free.back();
            g_pin_free.pop_back();
            return s;
        }
    }
    State *s = nullptr;
    if (cudaMallocHost(&s, sizeof(State)) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return s;
}
static void pinned_state_put(State *s) {
    std::lock_guard<std::mutex> lk(g_pin_mu);
    g_pin_free.push_back(s);
}

static void plan_free(cnmf_plan *p) {
    if (!p) return;
    for (cudaEvent_t e : p->tev) cudaEventDestroy(e);
    if (p->gexec) cudaGraphExecDestroy(p->gexec);
    if (p->scratch) cudaFree(p->scratch);
    if (p->h_state) pinned_state_put(p->h_state);
    if (p->ev_in) cudaEventDestroy(p->ev_in);
    if (p->ev_out) cudaEventDestroy(p->ev_out);
    if (p->own_stream) cudaStreamDestroy(p->own_stream);
    delete p;
}

}  // namespace cnmf
using namespace cnmf;

extern "C" int cnmf_version(void) { return CNMF_VERSION; }
extern "C" const char *cnmf_last_error(void) { return g_err; }
extern "C" int64_t cnmf_launch_count(void) { return g_launches.load(); }

extern "C" int64_t cnmf_padded_ld(int32_t m, int32_t dtype) {
    const int q = dtype == CNMF_F32 ? 4 : 2;  // elements per 16 bytes
    return ((int64_t)m + q - 1) / q * q;
}

extern "C" int cnmf_plan_create(cnmf_plan_t **plan, const cnmf_desc_t *desc) {
    if (!plan || !desc) {
        set_error("cnmf_plan_create: NULL argument");
        return CNMF_ERR_ARG;
    }
    *plan = nullptr;
    const cnmf_desc_t &d = *desc;
    if (d.dtype != CNMF_F32 && d.dtype != CNMF_F64) {
        set_error("dtype must be CNMF_F32 or CNMF_F64");
        return CNMF_ERR_ARG;
    }
    if (d.k < 1 || d.k > CNMF_MAX_K) {
        set_error("k=%d out of range 1..%d", d.k, CNMF_MAX_K);
        return CNMF_ERR_ARG;
    }
    if (d.m < 1 || d.n_rows < 1 || d.ld < d.m) {
        set_error("need n_rows >= 1, m >= 1, ld >= m (n_rows=%lld m=%d ld=%lld)", (long long)d.n_rows, d.m,
                  (long long)d.ld);
        return CNMF_ERR_ARG;
    }
    const int sz = d.dtype == CNMF_F32 ? 4 : 8;
    if ((d.ld * sz) % 16 != 0) {
        set_error("ld * sizeof(T) = %lld is not a multiple of 16 (use cnmf_padded_ld / cnmf_prepare)",
                  (long long)(d.ld * sz));
        return CNMF_ERR_ALIGN;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        set_error("no CUDA device: libcnmf_b200 has no CPU path");
        return CNMF_ERR_NODEVICE;
    }
    cnmf_plan *p = new cnmf_plan();
    p->d = d;
    if (p->d.eps <= 0) p->d.eps = default_eps(d.dtype);
    if (p->d.batch < 1) p->d.batch = 1;
    CU(cudaGetDevice(&p->device));
    // three attribute queries instead of cudaGetDeviceProperties (milliseconds per call, which a
    // 1 000 x 81 solve notices)
    int cc_major = 0, cc_minor = 0, n_sms = 0, max_smem = 0;
    CU(cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, p->device));
    CU(cudaDeviceGetAttribute(&cc_minor, cudaDevAttrComputeCapabilityMinor, p->device));
    CU(cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, p->device));
    CU(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, p->device));
    if (cc_major != 10) {
        set_error("device %d is sm_%d%d; this library is built for sm_100a only", p->device, cc_major, cc_minor);
        plan_free(p);
        return CNMF_ERR_NODEVICE;
    }
    p->num_sms = n_sms;
    p->use_graph = env_tune("GRAPH", 1) != 0;
    int rc = 0;
    if (p->d.batch > 1 || d.variant == CNMF_VARIANT_ONE_CTA || d.variant == CNMF_VARIANT_RESIDENT) {
        if (d.comm) {
            set_error("batched plans are not sharded: give each rank its own problems");
            plan_free(p);
            return CNMF_ERR_UNSUPPORTED;
        }
        p->batched = 1;
        p->batched_smem = batched_smem_bytes(d.dtype, d.k, d.ld, &p->batched_threads);
        {
            // problems whose X and V fit a cluster's shared memory stay resident there (CNMF_RESIDENT=0: off)
            // CNMF_RESIDENT: 0 = never, 1 = whenever it fits, unset = when all clusters of the batch are
            // co-resident (batch x CTAs per problem <= SMs).  With more problems than that the streaming
            // kernel wins: its two CTAs per SM hide each other's barriers (config 4: 0.42 vs 0.69 ms).
            const char *e = getenv("CNMF_RESIDENT");
            const bool never = e && e[0] == '0', always = (e && e[0] == '1') || d.variant == CNMF_VARIANT_RESIDENT;
            if (!never) {
                p->res_cs = resident_cluster_size(d.dtype, d.k, d.ld, d.n_rows, max_smem, &p->res_rows, &p->res_smem);
                const long long nb = p->d.batch > 1 ? p->d.batch : 1;
                if (!always && nb * p->res_cs > n_sms) p->res_cs = 0;
                // the device must be able to co-schedule one such cluster (it always can on an idle B200)
                if (p->res_cs > 0 && !resident_can_launch(d.dtype, d.k, d.ld, d.m, p->res_cs, p->res_rows)) p->res_cs = 0;
            }
        }
        if (p->res_cs > 0) {
            *plan = p;
            return 0;
        }
        if (d.variant == CNMF_VARIANT_RESIDENT) {
            set_error("resident plan: %lld x %d, k=%d does not fit a cluster's shared memory", (long long)d.n_rows, d.m, d.k);
            plan_free(p);
            return CNMF_ERR_UNSUPPORTED;
        }
        if (p->batched_smem <= 0 || p->batched_smem > max_smem) {
            set_error("batched mode: m=%d k=%d does not fit shared memory", d.m, d.k);
            plan_free(p);
            return CNMF_ERR_UNSUPPORTED;
        }
        *plan = p;
        return 0;
    }
    rc = choose_config(p->d, p->num_sms, max_smem, &p->cfg);
    if (rc) {
        plan_free(p);
        return rc;
    }
    p->n_tiles = (d.n_rows + p->cfg.R - 1) / p->cfg.R;
    p->n_entries = 2 * d.k * (int)d.ld;
    p->part_stride = p->n_entries;
    p->nblk = p->cfg.grid * p->cfg.S2 * (p->cfg.NT < 1 ? 1 : p->cfg.NT);
    p->n_chi = p->cfg.grid * p->cfg.chi_slots;
    cudaError_t e;
    const size_t part_b = (size_t)p->nblk * p->part_stride * sizeof(double);
    // one device allocation carved into [partial blocks | chi2 partials | reduced sums | state]
    auto up256 = [](size_t v) { return (v + 255) / 256 * 256; };
    const size_t chi_b = up256((size_t)p->n_chi * sizeof(double)), red_b = up256((size_t)(p->n_entries + 1) * sizeof(double));
    const size_t scratch_b = up256(part_b) + chi_b + red_b + up256(sizeof(State));
    e = cudaMalloc(&p->scratch, scratch_b);
    if (e == cudaSuccess) {
        char *base = (char *)p->scratch;
        p->part = (double *)base;
        p->chi_part = (double *)(base + up256(part_b));
        p->red = (double *)(base + up256(part_b) + chi_b);
        p->state = (State *)(base + up256(part_b) + chi_b + red_b);
        p->h_state = pinned_state_get();
        if (!p->h_state) e = cudaErrorMemoryAllocation;
    }
    if (e != cudaSuccess ||
        (e = cudaStreamCreateWithFlags(&p->own_stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&p->ev_in, cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&p->ev_out, cudaEventDisableTiming)) != cudaSuccess) {
        plan_free(p);
        return cuda_fail(e, "plan scratch allocation");
    }
    p->scratch_bytes = (long long)part_b + (p->n_chi + p->n_entries + 1) * 8LL + sizeof(State);
    CU(cudaMemset(p->state, 0, sizeof(State)));
    CU(cudaMemset(p->part, 0, part_b));  // blocks of the tail warp's extra row phases keep zeros outside its columns
    *plan = p;
    return 0;
}

extern "C" int cnmf_plan_destroy(cnmf_plan_t *plan) {
    plan_free(plan);
    return 0;
}

static void fill_info(const cnmf_desc_t &d, const StreamConfig &c, cnmf_info_t *info) {
    const int sz = d.dtype == CNMF_F32 ? 4 : 8;
    info->bytes_per_iter = (int64_t)d.n_rows * d.m * 2 * sz;
    info->grid = c.grid;
    info->threads = c.threads;
    info->smem_bytes = c.smem_bytes;
    info->stages = c.stages;
    info->tile_rows = c.R;
    info->lanes_per_row = c.G;
    info->h_streams = c.S2;
    info->partial_blocks = c.grid * c.S2 * (c.NT < 1 ? 1 : c.NT);
    info->variant = (c.wreg ? 100000 * c.crows : 0) + c.G * 1000 + c.C2 * 100 + c.RP1 * 10 + c.NW1;
    if (c.mma) info->variant = 900000 + c.mma * 1000 + c.crows * 100 + c.NW1;  // tensor-path kernels: form (5 = fp32), column parts, row-owner warps
    info->launches_per_iter = 3;
}

extern "C" int cnmf_plan_dryrun(const cnmf_desc_t *desc, cnmf_info_t *info) {
    if (!desc || !info) return CNMF_ERR_ARG;
    memset(info, 0, sizeof(*info));
    const cnmf_desc_t &d = *desc;
    if ((d.dtype != CNMF_F32 && d.dtype != CNMF_F64) || d.k < 1 || d.k > CNMF_MAX_K || d.m < 1 || d.n_rows < 1 ||
        d.ld < d.m) {
        set_error("cnmf_plan_dryrun: bad description");
        return CNMF_ERR_ARG;
    }
    if ((d.ld * (d.dtype == CNMF_F32 ? 4 : 8)) % 16) return CNMF_ERR_ALIGN;
    if (d.variant == CNMF_VARIANT_RESIDENT) {  // what a resident plan would be on a B200 (148 SMs, 227 KB)
        int rows = 0, smem = 0;
        const int cs = resident_cluster_size(d.dtype, d.k, d.ld, d.n_rows, 232448, &rows, &smem);
        if (cs <= 0) {
            set_error("resident plan: %lld x %d, k=%d does not fit a cluster's shared memory", (long long)d.n_rows, d.m, d.k);
            return CNMF_ERR_UNSUPPORTED;
        }
        const int nb = d.batch > 1 ? d.batch : 1;
        info->grid = nb * cs;
        info->threads = 512;
        info->smem_bytes = smem;
        info->variant = 700000 + cs;
        info->tile_rows = rows;
        info->bytes_per_iter = (int64_t)d.n_rows * d.m * 2 * (d.dtype == CNMF_F32 ? 4 : 8) * nb;
        return 0;
    }
    StreamConfig c;
    if (int rc = choose_config(d, 148, 232448, &c)) return rc;
    fill_info(d, c, info);
    return 0;
}

extern "C" int cnmf_plan_info(const cnmf_plan_t *p, cnmf_info_t *info) {
    if (!p || !info) return CNMF_ERR_ARG;
    memset(info, 0, sizeof(*info));
    const int sz = p->d.dtype == CNMF_F32 ? 4 : 8;
    info->bytes_per_iter = (int64_t)p->d.n_rows * p->d.m * 2 * sz * p->d.batch;
    if (p->batched) {
        info->grid = p->d.batch;
        info->threads = p->batched_threads;
        info->smem_bytes = p->batched_smem;
        info->launches_per_iter = 0;
        if (p->res_cs > 0) {  // resident cluster kernel: `variant` carries the cluster size, tile_rows the rows per CTA
            info->grid = p->d.batch * p->res_cs;
            info->threads = 512;
            info->smem_bytes = p->res_smem;
            info->variant = 700000 + p->res_cs;
            info->tile_rows = p->res_rows;
        }
        return 0;
    }
    fill_info(p->d, p->cfg, info);
    info->scratch_bytes = p->scratch_bytes;
    return 0;
}

// ---------------------------------------------------------------- iteration loop
namespace cnmf {

// The legacy default stream cannot be captured into a graph: work for a NULL
// stream runs on the plan's own stream, ordered after / before stream 0 by events.
struct StreamScope {
    cnmf_plan *p;
    cudaStream_t user, s;
    int begin(cnmf_plan *plan, void *stream) {
        p = plan;
        user = (cudaStream_t)stream;
        s = user;
        if (user == nullptr || user == cudaStreamLegacy) {
            CU(cudaEventRecord(p->ev_in, user));
            CU(cudaStreamWaitEvent(p->own_stream, p->ev_in, 0));
            s = p->own_stream;
        }
        return 0;
    }
    int end() {
        if (s != user) {
            CU(cudaEventRecord(p->ev_out, s));
            CU(cudaStreamWaitEvent(user, p->ev_out, 0));
        }
        return 0;
    }
};

static int enqueue_k1(cnmf_plan *p, const void *X, const void *V, void *W, const void *H, int update_w,
                      cudaStream_t st) {
    StreamArgs a;
    a.X = X;
    a.V = V;
    a.W = W;
    a.H = H;
    a.part = p->part;
    a.chi_part = p->chi_part;
    a.state = p->state;
    a.n_rows = p->d.n_rows;
    a.ld = p->d.ld;
    a.part_stride = p->part_stride;
    a.m = p->d.m;
    a.n_tiles = (int)p->n_tiles;
    a.update_w = update_w;
    a.eps = p->d.eps;
    int e = launch_stream(p->cfg, a, st);
    if (e) return cuda_fail((cudaError_t)e, "stream kernel launch");
    return 0;
}

// reduce partial blocks -> (allreduce over ranks) -> record chi2, test tolerance, update H
static int enqueue_rf(cnmf_plan *p, void *H, double *hist, double tol, int update_h, cudaStream_t st) {
    launch_reduce(p->part, p->nblk, p->part_stride, p->n_entries, p->chi_part, p->n_chi, p->red, p->state, st);
    CU(cudaGetLastError());
    if (comm_peer_ok(p->d.comm, (size_t)p->n_entries + 1)) {
        // all-reduce over peer memory fused with the finalise step: no NCCL on the iteration path
        launch_exchange(p->d.dtype, comm_peer_view(p->d.comm), p->red, p->n_entries + 1, H, p->state, hist, p->d.k, p->d.m,
                        p->d.ld, tol, p->d.eps, update_h, 1, st);
        CU(cudaGetLastError());
        return 0;
    }
    if (p->d.comm) {
        int e = comm_allreduce_f64(p->d.comm, p->red, (size_t)p->n_entries + 1, st);
        if (e) return e;
    }
    launch_finalize(p->d.dtype, p->red, H, p->state, hist, p->d.k, p->d.m, p->d.ld, tol, p->d.eps, update_h, st);
    CU(cudaGetLastError());
    return 0;
}

static int enqueue_iters(cnmf_plan *p, const void *X, const void *V, void *W, void *H, double *hist, double tol,
                         int update, int n, cudaStream_t st) {
    for (int i = 0; i < n; ++i) {
        if (int e = enqueue_rf(p, H, hist, tol, (update & CNMF_UPDATE_H) ? 1 : 0, st)) return e;
        if (p->timing) {
            if ((int)p->tev.size() < p->tev_used + 2) {
                cudaEvent_t a, b;
                CU(cudaEventCreate(&a));
                CU(cudaEventCreate(&b));
                p->tev.push_back(a);
                p->tev.push_back(b);
            }
            CU(cudaEventRecord(p->tev[p->tev_used], st));
        }
        if (int e = enqueue_k1(p, X, V, W, H, (update & CNMF_UPDATE_W) ? 1 : 0, st)) return e;
        if (p->timing) {
            CU(cudaEventRecord(p->tev[p->tev_used + 1], st));
            p->tev_used += 2;
        }
    }
    return 0;
}

int comm_sum_small(cnmf_comm_t *comm, double *d_buf, int count, void *d_state, cudaStream_t st) {
    if (!comm || comm->world_size <= 1) return 0;
    if (comm_peer_ok(comm, (size_t)count)) {
        launch_exchange(CNMF_F64, comm_peer_view(comm), d_buf, count, nullptr, (State *)d_state, nullptr, 1, 1, 2, 0.0, 0.0, 0,
                        0, st);
        CU(cudaGetLastError());
        State hs;
        CU(cudaMemcpyAsync(&hs, d_state, sizeof(State), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        if (hs.peer_timeout) {
            set_error("peer exchange timed out (a rank of the communicator is gone)");
            return CNMF_ERR_NCCL;
        }
        return 0;
    }
    return comm_allreduce_f64(comm, d_buf, (size_t)count, st);
}

// The graph of `n` iterations on these buffers (captured and instantiated once per plan and buffer set).
static int ensure_graph(cnmf_plan *p, const void *dX, const void *dV, void *dW, void *dH, double *d_hist, int n, int update,
                        double tol, cudaStream_t st) {
    GraphKey key{dX, dV, dW, dH, d_hist, n, update, tol, st};
    if (p->have_graph && key == p->gkey) return 0;
    if (p->gexec) {
        cudaGraphExecDestroy(p->gexec);
        p->gexec = nullptr;
        p->have_graph = false;
    }
    cudaGraph_t g = nullptr;
    CU(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    g_capturing = true;
    int e = enqueue_iters(p, dX, dV, dW, dH, d_hist, tol, update, n, st);
    g_capturing = false;
    cudaError_t ce = cudaStreamEndCapture(st, &g);
    if (e) {
        if (g) cudaGraphDestroy(g);
        return e;
    }
    if (ce != cudaSuccess) return cuda_fail(ce, "cudaStreamEndCapture");
    ce = cudaGraphInstantiate(&p->gexec, g, 0);
    cudaGraphDestroy(g);
    if (ce != cudaSuccess) return cuda_fail(ce, "cudaGraphInstantiate");
    // the first launch of an executable graph uploads it: do that here, off the iteration path
    ce = cudaGraphUpload(p->gexec, st);
    if (ce != cudaSuccess) return cuda_fail(ce, "cudaGraphUpload");
    p->gkey = key;
    p->have_graph = true;
    return 0;
}

static int check_ptr16(const void *q, const char *name) {
    if (!q) {
        set_error("%s is NULL", name);
        return CNMF_ERR_ARG;
    }
    if (((uintptr_t)q) & 15) {
        set_error("%s is not 16-byte aligned", name);
        return CNMF_ERR_ALIGN;
    }
    return 0;
}

}  // namespace cnmf

extern "C" int cnmf_solve(cnmf_plan_t *p, const void *dX, const void *dV, void *dW, void *dH, double *d_hist,
                          int32_t max_iter, double tol, int32_t check_every, int32_t update, int32_t *n_iter_out,
                          void *stream) {
    if (!p || p->batched) {
        set_error("cnmf_solve: NULL or batched plan");
        return CNMF_ERR_ARG;
    }
    if (int e = check_ptr16(dX, "X")) return e;
    if (int e = check_ptr16(dV, "V")) return e;
    if (int e = check_ptr16(dW, "W")) return e;
    const bool resume = (update & CNMF_SOLVE_RESUME) != 0;
    const bool time_kernels = (update & CNMF_SOLVE_TIME_KERNELS) != 0;
    const bool prepare_only = (update & CNMF_SOLVE_PREPARE) != 0;
    update &= ~(CNMF_SOLVE_RESUME | CNMF_SOLVE_TIME_KERNELS | CNMF_SOLVE_PREPARE);
    if (!dH || !d_hist || max_iter < 0 || (update & ~CNMF_UPDATE_BOTH)) {
        set_error("cnmf_solve: bad H / hist / max_iter / update");
        return CNMF_ERR_ARG;
    }
    StreamScope sc;
    if (int e = sc.begin(p, stream)) return e;
    cudaStream_t st = sc.s;
    if (prepare_only) {
        // plan-time work of a coming solve with these buffers: capture, instantiate and upload its graph
        const bool adaptive0 = tol > 0.0;
        const int chunk0 = adaptive0 ? (check_every > 0 ? check_every : 10) : 20;
        if (p->use_graph && max_iter >= chunk0)
            if (int e = ensure_graph(p, dX, dV, dW, dH, d_hist, chunk0, update, adaptive0 ? tol : 0.0, st)) return e;
        if (n_iter_out) *n_iter_out = 0;
        return sc.end();
    }
    p->timing = time_kernels;
    p->tev_used = 0;
    if (!resume) {
        CU(cudaMemsetAsync(p->state, 0, sizeof(State), st));
        if (int e = enqueue_k1(p, dX, dV, dW, dH, 0, st)) return e;  // partials + chi2 of (W0, H0)
    }
    const bool adaptive = tol > 0.0;
    if (!adaptive) tol = 0.0;
    int chunk = adaptive ? (check_every > 0 ? check_every : 10) : 20;
    int left = max_iter;
    bool stopped = false;
    while (left > 0 && !stopped) {
        const int n = left < chunk ? left : chunk;
        if (p->use_graph && !p->timing && n == chunk) {
            if (int e = ensure_graph(p, dX, dV, dW, dH, d_hist, n, update, tol, st)) return e;
            CU(cudaGraphLaunch(p->gexec, st));
            g_launches.fetch_add(3LL * n, std::memory_order_relaxed);  // reduce + finalize + fused kernel per iteration
        } else {
            if (int e = enqueue_iters(p, dX, dV, dW, dH, d_hist, tol, update, n, st)) return e;
        }
        left -= n;
        if (adaptive && left > 0) {
            CU(cudaMemcpyAsync(p->h_state, p->state, sizeof(State), cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            stopped = p->h_state->done != 0;
        }
    }
    if (int e = enqueue_rf(p, dH, d_hist, tol, 0, st)) return e;  // chi2 of the last pass
    // The device state is read back after EVERY solve (one 32-byte copy; the stream is drained here
    // anyway): a peer that never delivered its partial sums, or a chi2 that went non-finite, ends a
    // fixed-count solve with an error too -- after `done` is set the remaining kernels are no-ops and
    // the history behind n_iter is never written.
    int rc = 0;
    CU(cudaMemcpyAsync(p->h_state, p->state, sizeof(State), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (n_iter_out) *n_iter_out = p->h_state->done ? p->h_state->n_iter : max_iter;
    if (p->h_state->peer_timeout) {
        set_error("peer exchange timed out at iteration %d (a rank of the communicator is gone)", p->h_state->n_iter);
        rc = CNMF_ERR_NCCL;
    } else if (p->h_state->nonfinite) {
        set_error("chi2 became non-finite at iteration %d", p->h_state->n_iter);
        rc = CNMF_ERR_NONFINITE;
    }
    if (int e = sc.end()) return e;
    return rc;
}

// The sharded schedule with every rank on ONE device, driven from one host thread on one stream
// (include/cnmf_b200.h): per iteration each rank's reduce, ONE cooperative exchange launch for all
// ranks, each rank's fused kernel.
extern "C" int cnmf_solve_group(cnmf_plan_t **plans, int32_t world, const void *const *dX, const void *const *dV,
                                void *const *dW, void *const *dH, double *const *d_hist, int32_t max_iter, double tol,
                                int32_t update, int32_t *n_iter_out, void *stream) {
    if (!plans || world < 1 || world > kMaxPeers || !dX || !dV || !dW || !dH || !d_hist || max_iter < 0 ||
        (update & ~CNMF_UPDATE_BOTH)) {
        set_error("cnmf_solve_group: bad arguments");
        return CNMF_ERR_ARG;
    }
    int dev = -1;
    CU(cudaGetDevice(&dev));
    for (int r = 0; r < world; ++r) {
        cnmf_plan *p = plans[r];
        if (!p || p->batched || !p->d.comm || !p->d.comm->local || p->d.comm->world_size != world || p->d.comm->rank != r ||
            p->d.comm->device != dev || p->device != dev || p->d.k != plans[0]->d.k || p->d.m != plans[0]->d.m ||
            p->d.ld != plans[0]->d.ld || p->d.dtype != plans[0]->d.dtype || !comm_peer_ok(p->d.comm, (size_t)p->n_entries + 1)) {
            set_error("cnmf_solve_group: plan %d is not rank %d of a %d-rank local group on the current device %d", r, r, world, dev);
            return CNMF_ERR_ARG;
        }
        if (int e = check_ptr16(dX[r], "X")) return e;
        if (int e = check_ptr16(dV[r], "V")) return e;
        if (int e = check_ptr16(dW[r], "W")) return e;
        if (!dH[r] || !d_hist[r]) {
            set_error("cnmf_solve_group: NULL H / hist of rank %d", r);
            return CNMF_ERR_ARG;
        }
    }
    cnmf_plan *p0 = plans[0];
    cudaStream_t st = stream ? (cudaStream_t)stream : p0->own_stream;
    GroupRank h_ranks[kMaxPeers];
    for (int r = 0; r < world; ++r) {
        h_ranks[r].pv = comm_peer_view(plans[r]->d.comm);
        h_ranks[r].red = plans[r]->red;
        h_ranks[r].H = dH[r];
        h_ranks[r].state = plans[r]->state;
        h_ranks[r].hist = d_hist[r];
    }
    GroupRank *d_ranks = nullptr;
    CU(cudaMalloc((void **)&d_ranks, sizeof(GroupRank) * world));
    struct Free {
        void *q;
        ~Free() { cudaFree(q); }
    } free_ranks{d_ranks};
    CU(cudaMemcpyAsync(d_ranks, h_ranks, sizeof(GroupRank) * world, cudaMemcpyHostToDevice, st));
    if (!(tol > 0.0)) tol = 0.0;
    auto group_rf = [&](int update_h) -> int {
        for (int r = 0; r < world; ++r) {
            cnmf_plan *p = plans[r];
            launch_reduce(p->part, p->nblk, p->part_stride, p->n_entries, p->chi_part, p->n_chi, p->red, p->state, st);
        }
        CU(cudaGetLastError());
        int e = launch_exchange_group(p0->d.dtype, d_ranks, world, p0->n_entries + 1, p0->d.k, p0->d.m, p0->d.ld, tol, p0->d.eps,
                                      update_h, 1, st);
        if (e) return cuda_fail((cudaError_t)e, "cooperative exchange launch");
        return 0;
    };
    for (int r = 0; r < world; ++r) {
        CU(cudaMemsetAsync(plans[r]->state, 0, sizeof(State), st));
        if (int e = enqueue_k1(plans[r], dX[r], dV[r], dW[r], dH[r], 0, st)) return e;
    }
    bool stopped = false;
    for (int it = 0; it < max_iter && !stopped; ++it) {
        if (int e = group_rf((update & CNMF_UPDATE_H) ? 1 : 0)) return e;
        for (int r = 0; r < world; ++r)
            if (int e = enqueue_k1(plans[r], dX[r], dV[r], dW[r], dH[r], (update & CNMF_UPDATE_W) ? 1 : 0, st)) return e;
        if (tol > 0.0 && (it % 10) == 9) {
            CU(cudaMemcpyAsync(p0->h_state, p0->state, sizeof(State), cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            stopped = p0->h_state->done != 0;
        }
    }
    if (int e = group_rf(0)) return e;
    int rc = 0;
    for (int r = 0; r < world; ++r) {
        cnmf_plan *p = plans[r];
        CU(cudaMemcpyAsync(p->h_state, p->state, sizeof(State), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        if (n_iter_out) n_iter_out[r] = p->h_state->done ? p->h_state->n_iter : max_iter;
        if (p->h_state->peer_timeout) {
            set_error("peer exchange timed out at iteration %d (rank %d)", p->h_state->n_iter, r);
            rc = CNMF_ERR_NCCL;
        } else if (p->h_state->nonfinite && !rc) {
            set_error("chi2 became non-finite at iteration %d", p->h_state->n_iter);
            rc = CNMF_ERR_NONFINITE;
        }
    }
    return rc;
}

extern "C" int cnmf_plan_kernel_time(cnmf_plan_t *p, double *ms_total, int32_t *n_launches) {
    if (!p || !ms_total) return CNMF_ERR_ARG;
    double tot = 0.0;
    for (int i = 0; i + 1 < p->tev_used; i += 2) {
        CU(cudaEventSynchronize(p->tev[i + 1]));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, p->tev[i], p->tev[i + 1]));
        tot += ms;
    }
    *ms_total = tot;
    if (n_launches) *n_launches = p->tev_used / 2;
    return 0;
}

extern "C" int cnmf_chi2(cnmf_plan_t *p, const void *dX, const void *dV, const void *dW, const void *dH,
                         double *d_out, void *stream) {
    if (!p || p->batched || !d_out || !dH) {
        set_error("cnmf_chi2: bad arguments");
        return CNMF_ERR_ARG;
    }
    if (int e = check_ptr16(dX, "X")) return e;
    if (int e = check_ptr16(dV, "V")) return e;
    if (int e = check_ptr16(dW, "W")) return e;
    StreamScope sc;
    if (int e = sc.begin(p, stream)) return e;
    cudaStream_t st = sc.s;
    CU(cudaMemsetAsync(p->state, 0, sizeof(State), st));
    if (int e = enqueue_k1(p, dX, dV, const_cast<void *>(dW), dH, 0, st)) return e;
    launch_reduce(p->part, p->nblk, p->part_stride, p->n_entries, p->chi_part, p->n_chi, p->red, p->state, st);
    CU(cudaGetLastError());
    if (comm_peer_ok(p->d.comm, (size_t)p->n_entries + 1)) {
        launch_exchange(p->d.dtype, comm_peer_view(p->d.comm), p->red, p->n_entries + 1, nullptr, p->state, nullptr, p->d.k,
                        p->d.m, p->d.ld, 0.0, p->d.eps, 0, 0, st);
        CU(cudaGetLastError());
    } else if (p->d.comm)
        if (int e = comm_allreduce_f64(p->d.comm, p->red, (size_t)p->n_entries + 1, st)) return e;
    CU(cudaMemcpyAsync(d_out, p->red + p->n_entries, sizeof(double), cudaMemcpyDeviceToDevice, st));
    return sc.end();
}

extern "C" int cnmf_solve_batched(cnmf_plan_t *p, const void *dX, const void *dV, void *dW, void *dH,
                                  double *d_hist, int32_t *d_niter, int32_t max_iter, double tol, int32_t update,
                                  void *stream) {
    if (!p || !p->batched) {
        set_error("cnmf_solve_batched: plan was not created with batch > 1");
        return CNMF_ERR_ARG;
    }
    if (int e = check_ptr16(dX, "X")) return e;
    if (int e = check_ptr16(dV, "V")) return e;
    if (!dW || !dH || !d_hist || !d_niter || max_iter < 0) {
        set_error("cnmf_solve_batched: NULL buffer or negative max_iter");
        return CNMF_ERR_ARG;
    }
    BatchedArgs a;
    a.X = dX;
    a.V = dV;
    a.W = dW;
    a.H = dH;
    a.hist = d_hist;
    a.niter = d_niter;
    a.n_rows = p->d.n_rows;
    a.ld = p->d.ld;
    a.m = p->d.m;
    a.k = p->d.k;
    a.batch = p->d.batch;
    a.max_iter = max_iter;
    a.update = update;
    a.dtype = p->d.dtype;
    a.tol = tol > 0 ? tol : 0.0;
    a.eps = p->d.eps;
    int e = p->res_cs > 0 ? launch_resident(a, p->res_cs, p->res_rows, (cudaStream_t)stream)
                          : launch_batched(a, (cudaStream_t)stream);
    if (e) return cuda_fail((cudaError_t)e, "batched kernel launch");
    return 0;
}

extern "C" int cnmf_prepare(int32_t dtype, const void *src_x, const void *src_w, const uint8_t *src_mask,
                            int64_t src_ld, int64_t mask_ld, void *dst_x, void *dst_v, int64_t dst_ld,
                            int64_t n_rows, int32_t m, int32_t *d_status, void *stream) {
    if ((dtype != CNMF_F32 && dtype != CNMF_F64) || !src_x || !dst_x || !dst_v || n_rows < 0 || m < 1 ||
        src_ld < m || dst_ld < m || (src_mask && mask_ld < m)) {
        set_error("cnmf_prepare: bad arguments");
        return CNMF_ERR_ARG;
    }
    int e = launch_prepare(dtype, src_x, src_w, src_mask, src_ld, mask_ld, dst_x, dst_v, dst_ld, n_rows, m, d_status,
                           (cudaStream_t)stream);
    if (e) return cuda_fail((cudaError_t)e, "prepare kernel launch");
    return 0;
}

extern "C" int cnmf_synth(int32_t dtype, void *dX, void *dV, int64_t n_rows, int64_t row0, int32_t m, int64_t ld,
                          int32_t k, const double *d_hstar, uint64_t seed, int32_t vmode, double p_missing,
                          void *stream) {
    if ((dtype != CNMF_F32 && dtype != CNMF_F64) || !dX || !dV || !d_hstar || n_rows < 0 || m < 1 || ld < m ||
        k < 1 || k > CNMF_MAX_K || vmode < 0 || vmode > 3) {
        set_error("cnmf_synth: bad arguments");
        return CNMF_ERR_ARG;
    }
    int e = launch_synth(dtype, dX, dV, n_rows, row0, m, ld, k, d_hstar, seed, vmode, p_missing, (cudaStream_t)stream);
    if (e) return cuda_fail((cudaError_t)e, "synth kernel launch");
    return 0;
}
