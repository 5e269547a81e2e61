// Host-side launch entry points shared between the translation units of
// libcnmf_b200.so (not part of the public ABI; see include/cnmf_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "comm.h"

namespace cnmf {

struct State;

void count_launch();

constexpr int kMaxDevices = 64;  // per-device caches of function attributes

// K1 configuration chosen by the plan (api.cu) and consumed by the launcher.
struct StreamConfig {
    int dtype;   // 0 f32, 1 f64
    int k;
    int G;       // lanes per row in role W (8 or 16)
    int C2;      // packs per thread in role H (1 or 2)
    int RP1;     // rows per lane in role W (1 or 2)
    int R;       // rows per tile
    int NW1;     // role-W warps
    int NW2;     // role-H warps
    int S2;      // role-H streams
    int TPS;     // threads per stream
    int stages;
    int threads;
    int smem_bytes;
    int grid;
    int flush_every;
    int debug;   // CNMF_TUNE DBG= bits, timing probes only
    int prefetch;  // L2 prefetch distance in tiles (0 = off)
    int NT;        // copies of the last column-owner warp (row-interleaved), >= 1
    int chi_slots;  // chi2 partial slots per CTA
    int noprod;  // 1: no producer warp (the column-owner warp that frees a stage last refills it)
    int wreg;    // NI > 0: register-resident role W (G = 32): packs per lane, = ceil(packs per row / 32)
    int gate_ns, lookahead;  // tuning builds (StreamParams)
    int crows;   // wreg: rows of a role-W chunk; mma: column parts of a tile in role W (1, 2 or 4)
    int mma;     // tensor-path kernels: 0 off; fp64 (mma_kernel.cuh): 3 = 16 warps, 1 = 12 warps; fp32 (mma32_kernel.cuh): 5
    unsigned char role[32];  // role of each warp (stream_kernel.cuh StreamParams::role)
    unsigned char hg0[16], hgn[16];  // mma: first column group / number of groups of each column-owner warp
};

struct StreamArgs {
    const void *X;
    const void *V;
    void *W;
    const void *H;
    double *part;
    double *chi_part;
    State *state;
    long long n_rows, ld, part_stride;
    int m, n_tiles, update_w;
    double eps;
};

// CTA size cap the stream kernel is compiled for (sets its register budget)
constexpr int stream_max_threads(int k) { return k <= 6 ? 640 : (k <= 8 ? 512 : 384); }

// compiled instances of the register-resident role W: NI packs per lane need 2 * NI * k registers
// for H alone, so fp64 rows of 289..320 columns (NI = 10) stop at k = 6 and fp32 at k = 8
constexpr bool stream_wreg_ok(bool f32, int k, int ni) {
    return f32 ? ((ni == 3 || ni == 5) && k <= 8) : (ni == 10 && k <= 6);
}

// mma_kernel.cuh: column groups (of 8) a column-owner warp can keep accumulators for
constexpr int kMmaGroupsPerWarp = 10;

// mma32_kernel.cuh: 16-column groups a column-owner warp can keep accumulators for
constexpr int kMma32GroupsPerWarp = 3;

// returns cudaError_t as int; sets the max-dynamic-smem attribute on first use
int launch_stream(const StreamConfig &c, const StreamArgs &a, cudaStream_t st);
// false if (dtype, k, G, C2) has no compiled instance
bool stream_variant_exists(int dtype, int k, int G, int C2);

void launch_reduce(const double *part, int nblk, long long stride, int n_entries, const double *chi_part, int n_chi,
                   double *red, const State *state, cudaStream_t st);
void launch_finalize(int dtype, const double *red, void *H, State *state, double *hist, int k, int m, long long ld,
                     double tol, double eps, int update_h, cudaStream_t st);
// K2 for the row-sharded mode with peer windows attached: sum of `count` doubles over the ranks
// through peer memory and, with finalize != 0, the chi2 / tolerance / H update on the result
void launch_exchange(int dtype, const PeerView &pv, double *red, int count, void *H, State *state, double *hist, int k,
                     int m, long long ld, double tol, double eps, int update_h, int finalize, cudaStream_t st);

// one rank of a single-device group for launch_exchange_group (device array, one entry per CTA)
struct GroupRank {
    PeerView pv;
    double *red;
    void *H;
    State *state;
    double *hist;
};
int launch_exchange_group(int dtype, const GroupRank *d_ranks, int world, int count, int k, int m, long long ld, double tol,
                          double eps, int update_h, int finalize, cudaStream_t st);

int launch_prepare(int dtype, const void *src_x, const void *src_w, const uint8_t *src_mask, long long src_ld,
                   long long mask_ld, void *dst_x, void *dst_v, long long dst_ld, long long n_rows, int m,
                   int *d_status, cudaStream_t st);
int launch_synth(int dtype, void *dX, void *dV, long long n_rows, long long row0, int m, long long ld, int k,
                 const double *d_hstar, uint64_t seed, int vmode, double p_missing, cudaStream_t st);

struct BatchedArgs {
    const void *X;
    const void *V;
    void *W;
    void *H;
    double *hist;
    int *niter;
    long long n_rows, ld;
    int m, k, batch, max_iter, update, dtype;
    double tol, eps;
};
int launch_batched(const BatchedArgs &a, cudaStream_t st);
// resident.cu: one cluster of `cs` CTAs per problem, X and V kept in shared memory for all iterations.
// resident_cluster_size returns 0 when the shape does not fit (then the streaming batched kernel runs).
int resident_cluster_size(int dtype, int k, long long ld, long long n_rows, int max_smem, int *rows_per_out,
                          int *smem_out);
int launch_resident(const BatchedArgs &a, int cs, int rows_per, cudaStream_t st);
bool resident_can_launch(int dtype, int k, long long ld, int m, int cs, int rows_per);
int batched_smem_bytes(int dtype, int k, long long ld, int *threads_out);

}  // namespace cnmf
