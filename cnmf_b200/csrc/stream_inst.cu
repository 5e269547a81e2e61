// One translation unit per (dtype, k): compiled with -DCNMF_T=float|double
// -DCNMF_TN=f32|f64 -DCNMF_K=<k>.  Exposes launch_stream_<TN>_k<k>().
#include "launch.h"
#include "stream_kernel.cuh"

#ifndef CNMF_T
#error "compile with -DCNMF_T=float|double -DCNMF_TN=f32|f64 -DCNMF_K=k"
#endif

#define CNMF_CAT_(a, b, c) launch_stream_##a##_k##b
#define CNMF_CAT(a, b) CNMF_CAT_(a, b, )
#define CNMF_FN CNMF_CAT(CNMF_TN, CNMF_K)

namespace cnmf {

// register budget per thread follows from the CTA size cap (one CTA per SM)
constexpr int kMaxThreads = stream_max_threads(CNMF_K);

template <int G, int C2, int RP1, int MAXT, int NI = 0>
static int launch_one(const StreamConfig &c, const StreamArgs &a, cudaStream_t st) {
    typedef CNMF_T T;
    void (*kern)(const StreamParams<T>) = stream_kernel<T, CNMF_K, G, C2, RP1, MAXT, NI>;
    // the opt-in to more than 48 KB of dynamic shared memory is a per-device function attribute
    static int smem_set[kMaxDevices] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= kMaxDevices) return (int)cudaErrorInvalidDevice;
    if (c.smem_bytes > smem_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, c.smem_bytes);
        if (e != cudaSuccess) return (int)e;
        smem_set[dev] = c.smem_bytes;
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
    p.NW1 = c.NW1;
    for (int i = 0; i < 32; ++i) p.role[i] = c.role[i];
    p.NW2 = c.NW2;
    p.S2 = c.S2;
    p.TPS = c.TPS;
    p.stages = c.stages;
    p.stages_inv = (unsigned)((0x100000000ull + (unsigned)c.stages - 1) / (unsigned)c.stages);
    p.update_w = a.update_w;
    p.flush_every = c.flush_every;
    p.debug = c.debug;
    p.prefetch = c.prefetch;
    p.NT = c.NT < 1 ? 1 : c.NT;
    p.noprod = c.noprod;
    p.crows = c.crows;
    p.gate_ns = c.gate_ns;
    p.lookahead = c.lookahead;
    p.eps = (T)a.eps;
    if (c.threads > MAXT) return (int)cudaErrorInvalidConfiguration;
    kern<<<c.grid, c.threads, c.smem_bytes, st>>>(p);
    count_launch();
    return (int)cudaGetLastError();
}

// Compiled instances.  The product library carries what the plan's default configuration can
// select: 8 lanes per row, <= 384 threads (168 registers), C2 and RP1 in {1, 2}.  The tuning build
// (-DCNMF_TUNABLE, cnmf_b200/build.py --tune) adds the sweep variants CNMF_TUNE can ask for:
// 16 lanes per row, the 512 / 640-thread caps and the register-resident role W.
template <int G>
static int launch_g(const StreamConfig &c, const StreamArgs &a, cudaStream_t st) {
    const int C2 = c.C2, RP1 = c.RP1;
    if (c.threads <= 384) {
        if (C2 == 1 && RP1 == 1) return launch_one<G, 1, 1, 384>(c, a, st);
        if (C2 == 2 && RP1 == 1) return launch_one<G, 2, 1, 384>(c, a, st);
        if (C2 == 1 && RP1 == 2) return launch_one<G, 1, 2, 384>(c, a, st);
        if (C2 == 2 && RP1 == 2) return launch_one<G, 2, 2, 384>(c, a, st);
    }
#ifdef CNMF_TUNABLE
    if constexpr (kMaxThreads >= 512) {
        if (c.threads <= 512 && RP1 == 1) {
            if (C2 == 1) return launch_one<G, 1, 1, 512>(c, a, st);
            if (C2 == 2) return launch_one<G, 2, 1, 512>(c, a, st);
        }
    }
    if constexpr (kMaxThreads >= 640) {
        if (c.threads <= 640 && RP1 == 1 && C2 == 1) return launch_one<G, 1, 1, 640>(c, a, st);
    }
#endif
    return (int)cudaErrorInvalidConfiguration;
}

#ifdef CNMF_TUNABLE
// register-resident role W (stream_kernel NI > 0): the instances follow stream_wreg_ok()
template <int NI>
static int launch_wreg(const StreamConfig &c, const StreamArgs &a, cudaStream_t st) {
    if constexpr (stream_wreg_ok(sizeof(CNMF_T) == 4, CNMF_K, NI)) {
        if (c.threads > 384 || c.RP1 != 1) return (int)cudaErrorInvalidConfiguration;
        if (c.C2 == 1) return launch_one<32, 1, 1, 384, NI>(c, a, st);
        if (c.C2 == 2) return launch_one<32, 2, 1, 384, NI>(c, a, st);
    }
    return (int)cudaErrorInvalidConfiguration;
}
#endif

int CNMF_FN(const StreamConfig &c, const StreamArgs &a, cudaStream_t st) {
#ifdef CNMF_TUNABLE
    if (c.wreg == 3) return launch_wreg<3>(c, a, st);
    if (c.wreg == 5) return launch_wreg<5>(c, a, st);
    if (c.wreg == 10) return launch_wreg<10>(c, a, st);
    if (c.G == 16 && !c.wreg) return launch_g<16>(c, a, st);
#endif
    if (c.wreg) return (int)cudaErrorInvalidConfiguration;
    if (c.G == 8) return launch_g<8>(c, a, st);
    return (int)cudaErrorInvalidValue;
}

}  // namespace cnmf
