// One translation unit per k (fp32 tensor-path kernel, TF32 x 3): compiled with -DCNMF_K=<k>, k = 1..8.
// Exposes launch_mma32_k<k>().
#include <string.h>

#include "launch.h"
#include "mma32_kernel.cuh"

#ifndef CNMF_K
#error "compile with -DCNMF_K=k"
#endif

#define CNMF_TCAT_(b) launch_mma32_k##b
#define CNMF_TCAT(b) CNMF_TCAT_(b)
#define CNMF_TFN CNMF_TCAT(CNMF_K)

namespace cnmf {

#ifdef CNMF_TUNABLE
template <int GHMAX, int MAXT>
static int launch_mma32_one(const StreamConfig &c, const StreamArgs &a, cudaStream_t st) {
    void (*kern)(const StreamParams<float>) = mma32_kernel<CNMF_K, GHMAX, MAXT>;
    static int smem_set[kMaxDevices] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= kMaxDevices) return (int)cudaErrorInvalidDevice;
    if (c.smem_bytes > smem_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, c.smem_bytes);
        if (e != cudaSuccess) return (int)e;
        smem_set[dev] = c.smem_bytes;
    }
    StreamParams<float> p;
    memset(&p, 0, sizeof(p));
    p.X = (const float *)a.X;
    p.V = (const float *)a.V;
    p.W = (float *)a.W;
    p.H = (const float *)a.H;
    p.part = a.part;
    p.chi_part = a.chi_part;
    p.state = a.state;
    p.n_rows = a.n_rows;
    p.ld = a.ld;
    p.part_stride = a.part_stride;
    p.m = a.m;
    p.n_tiles = a.n_tiles;
    p.R = 16;
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
    p.flush_every = c.flush_every;
    p.eps = (float)a.eps;
    if (c.threads > MAXT) return (int)cudaErrorInvalidConfiguration;
    kern<<<c.grid, c.threads, c.smem_bytes, st>>>(p);
    count_launch();
    return (int)cudaGetLastError();
}
#endif

int CNMF_TFN(const StreamConfig &c, const StreamArgs &a, cudaStream_t st) {
#ifdef CNMF_TUNABLE
    // 16 warps: 8 column owners of up to three 16-column groups, 7 row owners, the producer (20 warps measured 3 % slower)
    return launch_mma32_one<kMma32GroupsPerWarp, 512>(c, a, st);
#else
    // measured slower than the FFMA2 stream kernel: the product library does not carry it
    (void)c; (void)a; (void)st;
    return (int)cudaErrorInvalidConfiguration;
#endif
}

}  // namespace cnmf
