// One translation unit per (dtype, k): compiled with -DCNMF_T=float|double
// -DCNMF_TN=f32|f64 -DCNMF_K=<k>.  Exposes launch_stream_<TN>_k<k>().
#include "stream_kernel.cuh"
#include "launch.h"

#ifndef CNMF_T
#error "compile with -DCNMF_T=float|double -DCNMF_TN=f32|f64 -DCNMF_K=k"
#endif

#define CNMF_CAT_(a, b, c) launch_stream_##a##_k##b
#define CNMF_CAT(a, b) CNMF_CAT_(a, b, )
#define CNMF_FN CNMF_CAT(CNMF_TN, CNMF_K)

namespace cnmf {

// register budget per thread follows from the CTA size cap (one CTA per SM)
constexpr int kMaxThreads = stream_max_threads(CNMF_K);

template <int G, int C2>
static int launch_one(const StreamConfig &c, const StreamArgs &a, cudaStream_t st) {
    typedef CNMF_T T;
    auto kern = stream_kernel<T, CNMF_K, G, C2, kMaxThreads>;
    static int smem_set = 0;
    if (c.smem_bytes > smem_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, c.smem_bytes);
        if (e != cudaSuccess) return (int)e;
        smem_set = c.smem_bytes;
    }
    StreamParams<T> p;
    p.X = (const T *)a.X;
    p.V = (const T *)a.V;
    p.W = (T *)a.W;
    p.H = (const T *)a.H;
    p.part = a.part;
    p.chi_part = a.chi_part;
    p.state = a.state;
    p.n_rows = a.n_rows;
    p.ld = a.ld;
    p.part_stride = a.part_stride;
    p.m = a.m;
    p.n_tiles = a.n_tiles;
    p.R = c.R;
    p.NG1 = c.NG1;
    p.NW2 = c.NW2;
    p.S2 = c.S2;
    p.TPS = c.TPS;
    p.stages = c.stages;
    p.update_w = a.update_w;
    p.flush_every = c.flush_every;
    p.eps = (T)a.eps;
    if (c.threads > kMaxThreads) return (int)cudaErrorInvalidConfiguration;
    kern<<<c.grid, c.threads, c.smem_bytes, st>>>(p);
    count_launch();
    return (int)cudaGetLastError();
}

int CNMF_FN(const StreamConfig &c, const StreamArgs &a, cudaStream_t st) {
    if (c.G == 8 && c.C2 == 1) return launch_one<8, 1>(c, a, st);
    if (c.G == 8 && c.C2 == 2) return launch_one<8, 2>(c, a, st);
    if (c.G == 16 && c.C2 == 1) return launch_one<16, 1>(c, a, st);
    if (c.G == 16 && c.C2 == 2) return launch_one<16, 2>(c, a, st);
    return (int)cudaErrorInvalidValue;
}

}  // namespace cnmf
