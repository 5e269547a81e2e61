// One translation unit per k (fp64 only): compiled with -DCNMF_K=<k>, k = 1..8.  Exposes launch_mma_k<k>().
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "launch.h"
#include "mma_kernel.cuh"

#ifndef CNMF_K
#error "compile with -DCNMF_K=k"
#endif

#define CNMF_MCAT_(b) launch_mma_k##b
#define CNMF_MCAT(b) CNMF_MCAT_(b)
#define CNMF_MFN CNMF_MCAT(CNMF_K)

namespace cnmf {

template <int GHMAX, int MAXT>
static int launch_mma_one(const StreamConfig &c, const StreamArgs &a, cudaStream_t st) {
    void (*kern)(const StreamParams<double>) = mma_kernel<CNMF_K, GHMAX, MAXT>;
    static int smem_set[kMaxDevices] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= kMaxDevices) return (int)cudaErrorInvalidDevice;
    if (c.smem_bytes > smem_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, c.smem_bytes);
        if (e != cudaSuccess) return (int)e;
        smem_set[dev] = c.smem_bytes;
    }
    StreamParams<double> p;
    memset(&p, 0, sizeof(p));
    p.X = (const double *)a.X;
    p.V = (const double *)a.V;
    p.W = (double *)a.W;
    p.H = (const double *)a.H;
    p.part = a.part;
    p.chi_part = a.chi_part;
    p.state = a.state;
    p.n_rows = a.n_rows;
    p.ld = a.ld;
    p.part_stride = a.part_stride;
    p.m = a.m;
    p.n_tiles = a.n_tiles;
    p.R = 8;
    p.NW1 = c.NW1;
    p.NW2 = c.NW2;
    for (int i = 0; i < 32; ++i) p.role[i] = c.role[i];
    for (int i = 0; i < 16; ++i) {
        p.hg0[i] = c.hg0[i];
        p.hgn[i] = c.hgn[i];
    }
    p.stages = c.stages;
    p.stages_inv = (unsigned)((0x100000000ull + (unsigned)c.stages - 1) / (unsigned)c.stages);
    p.update_w = a.update_w;
    p.crows = c.crows;  // column parts of a tile in role W
    p.gate_ns = c.gate_ns;
    p.prefetch = c.prefetch;
    p.lookahead = c.lookahead;  // suspend-time hint of the mbarrier waits, ns
    p.eps = a.eps;
    if (c.threads > MAXT) return (int)cudaErrorInvalidConfiguration;
#ifdef CNMF_PROF
    // diagnostics build: 8 words behind the layout hold the issue time of each stage's copies
    if (c.smem_bytes + 64 <= 232448) {
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, c.smem_bytes + 64);
        const bool trace = getenv("CNMF_TRACE") != nullptr && a.update_w;
        static unsigned long long h_tr[CNMF_TRACE_TILES][6];
        if (trace) {
            for (int i = 0; i < CNMF_TRACE_TILES; ++i) {
                h_tr[i][0] = h_tr[i][2] = h_tr[i][3] = h_tr[i][5] = 0ull;
                h_tr[i][1] = h_tr[i][4] = ~0ull;
            }
            cudaMemcpyToSymbol(g_mma_trace, h_tr, sizeof(h_tr));
        }
        kern<<<c.grid, c.threads, c.smem_bytes + 64, st>>>(p);
        count_launch();
        if (trace) {
            cudaStreamSynchronize(st);
            cudaMemcpyFromSymbol(h_tr, g_mma_trace, sizeof(h_tr));
            const unsigned long long t0 = h_tr[300][0];
            printf("trace of CTA 0, tiles 300..340 (cycles from the issue of tile 300): issued, landed (first row owner), last part out of its loop, W' handed over, first column owner in, stage free\n");
            for (int i = 300; i < 340; ++i)
                printf("tile %3d stage %d: %7lld %7lld %7lld %7lld %7lld %7lld | copy %5lld  W pass %5lld  tail %5lld  queue %5lld  H pass %5lld\n", i, i % c.stages,
                       (long long)(h_tr[i][0] - t0), (long long)(h_tr[i][1] - t0), (long long)(h_tr[i][2] - t0), (long long)(h_tr[i][3] - t0),
                       (long long)(h_tr[i][4] - t0), (long long)(h_tr[i][5] - t0), (long long)(h_tr[i][1] - h_tr[i][0]), (long long)(h_tr[i][2] - h_tr[i][1]),
                       (long long)(h_tr[i][3] - h_tr[i][2]), (long long)(h_tr[i][4] - h_tr[i][3]), (long long)(h_tr[i][5] - h_tr[i][4]));
        }
        return (int)cudaGetLastError();
    }
#endif
    kern<<<c.grid, c.threads, c.smem_bytes, st>>>(p);
    count_launch();
    return (int)cudaGetLastError();
}

int CNMF_MFN(const StreamConfig &c, const StreamArgs &a, cudaStream_t st) {
    // c.mma: 3 = eight column-owner warps of up to five column groups each (16 warps, the default form),
    //        1 = four column-owner warps of up to ten groups (12 warps)
    if (c.mma == 3) return launch_mma_one<kMmaGroupsPerWarp / 2 + 1, 512>(c, a, st);
    if (c.mma == 1) return launch_mma_one<kMmaGroupsPerWarp, 384>(c, a, st);
#ifdef CNMF_TUNABLE
    if (c.mma == 4) return launch_mma_one<kMmaGroupsPerWarp / 2 + 1, 544>(c, a, st);  // 8 + 8 + producer (tuning builds)
#endif
    return (int)cudaErrorInvalidConfiguration;
}

}  // namespace cnmf
