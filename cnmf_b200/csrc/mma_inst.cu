// One translation unit per k (fp64 only): compiled with -DCNMF_K=<k>, k = 1..8.  Exposes launch_mma_k<k>().
#include "launch.h"
#include "mma_kernel.cuh"

#ifndef CNMF_K
#error "compile with -DCNMF_K=k"
#endif

#define CNMF_MCAT_(b) launch_mma_k##b
#define CNMF_MCAT(b) CNMF_MCAT_(b)
#define CNMF_MFN CNMF_MCAT(CNMF_K)

namespace cnmf {

template <int GHMAX, int MAXT, bool HREG>
static int launch_mma_one(const StreamConfig &c, const StreamArgs &a, cudaStream_t st) {
    void (*kern)(const StreamParams<double>) = mma_kernel<CNMF_K, GHMAX, MAXT, HREG>;
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
    p.stages = c.stages;
    p.stages_inv = (unsigned)((0x100000000ull + (unsigned)c.stages - 1) / (unsigned)c.stages);
    p.update_w = a.update_w;
    p.crows = c.crows;  // column parts of a tile in role W
    p.gate_ns = c.gate_ns;
    p.prefetch = c.prefetch;
    p.eps = a.eps;
    if (c.threads > MAXT) return (int)cudaErrorInvalidConfiguration;
#ifdef CNMF_PROF
    // diagnostics build: 8 words behind the layout hold the issue time of each stage's copies
    if (c.smem_bytes + 64 <= 232448) {
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, c.smem_bytes + 64);
        kern<<<c.grid, c.threads, c.smem_bytes + 64, st>>>(p);
        count_launch();
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
    if (c.mma == 3) return launch_mma_one<kMmaGroupsPerWarp / 2, 512, false>(c, a, st);
    if (c.mma == 1) return launch_mma_one<kMmaGroupsPerWarp, 384, false>(c, a, st);
    return (int)cudaErrorInvalidConfiguration;
}

}  // namespace cnmf
